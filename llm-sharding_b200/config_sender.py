"""``ConfigSender``: same wire format and API as the reference's control plane
(utils/config_sender.py:4-47) — a 6-key JSON pushed over ZeroMQ to ``tcp://ip:40700+k``.  Kept
so that existing launch scripts (send_config.py) drive B200 nodes unchanged; it carries no
compute and is not accelerated."""
import zmq

CONFIG_KEYS = ("src_addr", "dst_addr", "can_receive_user_request", "first_node_addr", "shards_start", "shards_end")


class ConfigSender:
    def __init__(self, node_port: int = 40700):
        self.node_port = node_port
        self.node_addr = ""
        self.config = dict()
        self._ctx = zmq.Context()
        self.send_socket = self._ctx.socket(zmq.PUSH)

    def build_config(self, shards_start: int, shards_end: int, can_receive_user_request: bool,
                     src_addr: str, dst_addr: str, first_node_addr: str = ""):
        if can_receive_user_request and first_node_addr == "":
            raise ValueError("first_node_addr cannot be empty when can_receive_user_request = True")
        values = (src_addr, dst_addr, can_receive_user_request, first_node_addr, shards_start, shards_end)
        self.config = dict(zip(CONFIG_KEYS, values))

    def send_config(self, node_ip: str) -> None:
        if self.node_addr != "":
            self.send_socket.disconnect(self.node_addr)
        self.node_addr = f"tcp://{node_ip}:{self.node_port}"
        self.send_socket.connect(self.node_addr)
        self.send_socket.send_json(self.config)

    def close(self) -> None:
        self.send_socket.close(0)
        self._ctx.term()

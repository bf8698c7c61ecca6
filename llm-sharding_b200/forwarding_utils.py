"""``build_position_ids`` with the reference's signature (utils/forwarding_utils.py:4-26).  On the
B200 path the positions collapse to one integer per request (the cache length) that the kernels
read on the device; this helper stays for callers that still want the [B,S] tensor."""
import torch


def build_position_ids(past_key_value, seq_len, device, batch_size: int = 1):
    if past_key_value is None:
        past_len = 0
    elif hasattr(past_key_value, "get_seq_length"):
        past_len = int(past_key_value.get_seq_length())
    elif isinstance(past_key_value, tuple) and len(past_key_value) == 2:
        past_len = int(past_key_value[0].shape[-2])
    else:
        raise ValueError("[ERROR] Unsupported past_key_value structure")
    pos = torch.arange(past_len, past_len + seq_len, device=device, dtype=torch.long)
    return pos.unsqueeze(0).expand(batch_size, -1).contiguous()

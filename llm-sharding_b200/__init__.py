"""b200-shard-chain: the shard-execution hot path of seanbonjean/llm-sharding
(``NodeWorker.pass_through_shard``) on hand-written sm_100a CUDA kernels behind a C ABI
(``include/b200shard.h``, ``lib/libb200shard.so``).

Host side mirrors the reference's operator interface (same class / method names, argument
meaning and error behaviour): ``LlamaShardPart`` (utils/shard_loader.py), ``NodeWorker`` /
``NodeController`` / ``Communicator`` (utils/node_worker.py), ``ConfigSender``
(utils/config_sender.py), ``build_position_ids`` (utils/forwarding_utils.py).

There is no CPU path: anything that computes raises if the CUDA library or a GPU is missing.
"""
__version__ = "0.1.0"

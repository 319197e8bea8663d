"""Data sources: on-device synthetic shards and the file-backed Herbarium path."""
from .dataset import GetData, create_dataset, load_index, shard_rows  # noqa: F401
from .synthetic import HostStagedShard, SyntheticShard  # noqa: F401

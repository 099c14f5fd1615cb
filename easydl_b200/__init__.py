"""easydl_b200: the B200-native sparse-embedding step of an EasyDL ElasticTrainer PS job.

Hot path only: id dedup, hash-mod shard partition, embedding gather,
segmented gradient reduce, Adagrad/Adam row update, elastic re-shard -- CUDA (sm_100a) behind
the C ABI in include/edlsparse.h, with a Python host layer that keeps the parameter-server /
worker names of the ElasticDL lineage.
"""
__version__ = "0.1.0"

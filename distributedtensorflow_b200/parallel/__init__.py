"""Parameter-server transports: in-process, TCP (CPU path), NVLink peer memory (B200 product
path) and the NCCL comparator; plus the shard state machine and variable layout."""

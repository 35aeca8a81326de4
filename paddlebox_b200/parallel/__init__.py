"""Multi-GPU: sharded sparse engine over NVLink peer memory, symmetric-memory dense collectives,
dense sync modes (per-step allreduce, K-step averaging, ZeRO-1 sharding, async host table)."""

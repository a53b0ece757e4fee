"""Benchmarks: allreduce latency/bandwidth sweep vs NCCL, scaling runs, roofline reports."""

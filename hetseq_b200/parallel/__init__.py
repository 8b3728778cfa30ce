"""Data parallelism: process-group plumbing, flat buffers, NCCL baseline and fused NVLink backend."""

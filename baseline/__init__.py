"""Reconstruction baseline: the reference's train steps rebuilt from stock
components only - PyTorch eager ops (cuBLAS GEMMs, ATen element-wise kernels, SDPA attention), NCCL
collectives through torch.distributed, torch's fused AdamW. None of this repository's kernels or
engines are used. It stands in for the reference's own JAX/XLA:GPU build, which cannot be installed
on this image ("Reference arm")."""

"""One argparse factory for all recipes.

Flags and defaults are exactly the reference's (main-single.py:155-167,
main-ddp.py:191-203, main-fsdp.py:206-219, main-pipe.py:224-236); additive flags are
grouped under "b200 extras" and never change the defaults' behaviour.
"""
from __future__ import annotations

import argparse


def build_parser(recipe: str = "single") -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description=f"b200-distributed-gpt recipe: {recipe}")
    parser.add_argument("--batch_size", type=int, default=64)
    parser.add_argument("--epochs", type=int, default=5)
    parser.add_argument("--sequence_length", type=int, default=256)
    parser.add_argument("--dim", type=int, default=256)
    parser.add_argument("--head_dim", type=int, default=32)
    parser.add_argument("--heads", type=int, default=8)
    parser.add_argument("--num_layers", type=int, default=8)
    parser.add_argument("--learning_rate", type=float, default=1e-4)
    parser.add_argument("--dataset_slice", type=str, default="100%")
    parser.add_argument("--num_workers", type=int, default=4)
    parser.add_argument("--disable_amp", action="store_true")
    parser.add_argument("--disable_compile", action="store_true",
                        help="accepted for CLI parity; this framework has no tracing compiler "
                             "(CUDA graphs are controlled by --disable_cuda_graph)")
    if recipe == "fsdp":
        parser.add_argument("--cpu_offload", action="store_true")
    x = parser.add_argument_group("b200 extras (additive)")
    x.add_argument("--synthetic", action="store_true",
                   help="synthetic token stream instead of TinyStories (no network needed)")
    x.add_argument("--synthetic_samples", type=int, default=4096)
    x.add_argument("--vocab_size", type=int, default=50257)
    x.add_argument("--max_steps", type=int, default=0, help="stop each epoch after N steps (0 = all)")
    x.add_argument("--seed", type=int, default=0)
    x.add_argument("--resume", type=str, default=None,
                   help="checkpoint to load before training; the optimizer state saved next to it "
                        "(<file>.resume.rank<r>) is restored when the world layout matches")
    x.add_argument("--resume_epochs", action="store_true",
                   help="with --resume: also skip the epochs the checkpoint had already completed")
    x.add_argument("--checkpoint_dir", type=str, default="checkpoints")
    x.add_argument("--no_save", action="store_true")
    x.add_argument("--no_generate", action="store_true")
    x.add_argument("--metrics_jsonl", type=str, default=None)
    x.add_argument("--disable_cuda_graph", action="store_true")
    x.add_argument("--profile", action="store_true",
                   help="NVTX ranges around train_step / forward / backward / grad-sync / optimizer "
                        "(runs the step eagerly instead of as one CUDA graph so the ranges bracket kernels)")
    if recipe in ("ddp", "pipe-ddp"):
        x.add_argument("--bucket_mb", type=float, default=16.0,
                       help="gradient bucket size of the DDP reducer (16 MiB: five decoder layers of the default GPT per all-reduce)")
        x.add_argument("--comm", choices=["auto", "symm", "nccl"], default="auto",
                       help="gradient all-reduce transport: own NVLink kernels or NCCL")
        x.add_argument("--comm_dtype", choices=["fp32", "bf16"], default="fp32",
                       help="own NVLink all-reduce: what crosses the wire (bf16 = half the bytes, fp32 accumulation "
                            "in the switch, one extra bf16 rounding of every addend and of the sum)")
    if recipe == "fsdp":
        x.add_argument("--comm", choices=["auto", "symm", "nccl"], default="auto")
        x.add_argument("--reshard_after_forward", action="store_true",
                       help="memory-sharded mode (the reference's FULL_SHARD): gathered weights and gradients live in "
                            "two rotating unit-sized buffers, units are re-gathered for their backward")
    if recipe in ("pipe", "pipe-ddp"):
        x.add_argument("--num_stages", type=int, default=0, help="0 = world size (pipe) / 2 (pipe-ddp)")
        x.add_argument("--num_microbatches", type=int, default=0, help="0 = num_stages (reference chunks=num_stages)")
        x.add_argument("--schedule", choices=["gpipe", "1f1b"], default="gpipe")
        x.add_argument("--checkpoint", choices=["never", "except_last", "always"], default="never",
                       help="activation checkpointing per micro-batch: re-run its forward inside its backward "
                            "(legacy Pipe's default was except_last, reference main-pipe.py:79-83)")
    return parser

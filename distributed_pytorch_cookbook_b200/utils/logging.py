"""Console + JSONL metrics and device-side step timing (the reference
only prints and uses tqdm; we keep its console strings and add structured output)."""
from __future__ import annotations

import json
import logging
import sys
import time
from pathlib import Path
from typing import Optional

import torch


def get_logger(name: str = "b200gpt", rank: int = 0) -> logging.Logger:
    logger = logging.getLogger(name)
    if not logger.handlers:
        h = logging.StreamHandler(sys.stdout)
        h.setFormatter(logging.Formatter(f"[%(asctime)s r{rank}] %(message)s", "%H:%M:%S"))
        logger.addHandler(h)
        logger.setLevel(logging.INFO if rank == 0 else logging.WARNING)
        logger.propagate = False
    return logger


class JsonlLogger:
    """Append-only JSONL metrics file (step time, tokens/s, loss, exposed comm ...)."""

    def __init__(self, path: Optional[str], enabled: bool = True):
        self.f = None
        if path and enabled:
            Path(path).parent.mkdir(parents=True, exist_ok=True)
            self.f = open(path, "a")

    def log(self, **kv):
        if self.f is not None:
            kv.setdefault("ts", time.time())
            self.f.write(json.dumps(kv) + "\n")
            self.f.flush()

    def close(self):
        if self.f is not None:
            self.f.close()
            self.f = None


class StepTimer:
    """CUDA-event (or perf_counter on CPU) timer around a region on the current stream."""

    def __init__(self, device):
        self.cuda = torch.device(device).type == "cuda"
        if self.cuda:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
        self.t0 = 0.0

    def start(self):
        if self.cuda:
            self.e0.record()
        else:
            self.t0 = time.perf_counter()

    def stop_ms(self) -> float:
        if self.cuda:
            self.e1.record()
            self.e1.synchronize()
            return self.e0.elapsed_time(self.e1)
        return (time.perf_counter() - self.t0) * 1e3

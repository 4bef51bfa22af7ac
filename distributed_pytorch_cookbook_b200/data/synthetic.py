"""Synthetic token stream (no network): the shape/dtype contract of the reference's
tokenised dataset (data.py:23-36: every item ``{input_ids[int64,S], attention_mask[int64,S]}``)
with ids uniform in ``[0, vocab)`` - SURVEY 2.8 R8.  Also a byte-level stand-in tokenizer so
``generate`` works offline, and a pinned-memory double-buffered prefetcher that overlaps
the H2D copy of batch i+1 with step i on a side stream (reference K13: its copies are
effectively synchronous)."""
from __future__ import annotations

from typing import Dict, Iterator, Optional

import torch
from torch.utils.data import DataLoader, Dataset


class SyntheticTokenDataset(Dataset):
    def __init__(self, num_samples: int, sequence_length: int, vocab_size: int = 50257,
                 seed: int = 0, pad_fraction: float = 0.0, pad_id: int = 2):
        g = torch.Generator().manual_seed(seed)
        self.ids = torch.randint(0, vocab_size, (num_samples, sequence_length), generator=g, dtype=torch.int64)
        self.mask = torch.ones_like(self.ids)
        if pad_fraction > 0:  # right-padding like tokenizer(padding="max_length")
            lens = torch.randint(int(sequence_length * (1 - pad_fraction)), sequence_length + 1,
                                 (num_samples,), generator=g)
            ar = torch.arange(sequence_length).unsqueeze(0)
            self.mask = (ar < lens.unsqueeze(1)).to(torch.int64)
            self.ids = torch.where(self.mask.bool(), self.ids, torch.full_like(self.ids, pad_id))

    def __len__(self):
        return self.ids.shape[0]

    def __getitem__(self, i) -> Dict[str, torch.Tensor]:
        return {"input_ids": self.ids[i], "attention_mask": self.mask[i]}


class SyntheticTokenizer:
    """Byte-level stand-in for GPT2Tokenizer (offline): ids 0..255 are bytes."""

    def __init__(self, vocab_size: int = 50257, model_max_length: int = 512):
        self.vocab_size = vocab_size
        self.model_max_length = model_max_length
        self.pad_token_id = 2
        self.eos_token_id = vocab_size - 1

    def __call__(self, texts, truncation=True, max_length=None, return_tensors=None, padding=False):
        if isinstance(texts, str):
            texts = [texts]
        rows = [list(t.encode("utf-8"))[: (max_length or self.model_max_length)] for t in texts]
        L = max(len(r) for r in rows) if padding != "max_length" else (max_length or self.model_max_length)
        ids = torch.full((len(rows), L), self.pad_token_id, dtype=torch.int64)
        am = torch.zeros((len(rows), L), dtype=torch.int64)
        for i, r in enumerate(rows):
            ids[i, : len(r)] = torch.tensor(r, dtype=torch.int64)
            am[i, : len(r)] = 1
        return {"input_ids": ids, "attention_mask": am}

    def decode(self, ids, skip_special_tokens=True):
        b = bytes(int(i) for i in ids if 0 <= int(i) < 256)
        return b.decode("utf-8", errors="replace")


def make_loader(dataset, batch_size: int, sampler=None, shuffle: bool = False, num_workers: int = 0,
                pin_memory: bool = None, drop_last: bool = True) -> DataLoader:
    if pin_memory is None:
        pin_memory = torch.cuda.is_available()
    return DataLoader(dataset, batch_size=batch_size, sampler=sampler,
                      shuffle=shuffle if sampler is None else False, num_workers=num_workers,
                      pin_memory=pin_memory, drop_last=drop_last)


class PinnedPrefetcher:
    """Wrap a loader of ``{input_ids, attention_mask}`` batches: stage each batch in a
    pinned buffer and copy it to ``device`` on a side stream one step ahead."""

    def __init__(self, loader, device, depth: int = 2):
        self.loader, self.device, self.depth = loader, torch.device(device), depth
        self.cuda = self.device.type == "cuda"
        self.stream = torch.cuda.Stream(self.device) if self.cuda else None

    def __len__(self):
        return len(self.loader)

    def _stage(self, batch):
        if not self.cuda:
            return batch, None
        with torch.cuda.stream(self.stream):
            out = {k: (v if v.is_pinned() else v.pin_memory()).to(self.device, non_blocking=True)
                   for k, v in batch.items()}
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return out, ev

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        it = iter(self.loader)
        queue = []
        for _ in range(self.depth):
            b = next(it, None)
            if b is None:
                break
            queue.append(self._stage(b))
        while queue:
            out, ev = queue.pop(0)
            if ev is not None:
                torch.cuda.current_stream(self.device).wait_event(ev)
                for v in out.values():
                    v.record_stream(torch.cuda.current_stream(self.device))
            b = next(it, None)
            if b is not None:
                queue.append(self._stage(b))
            yield out

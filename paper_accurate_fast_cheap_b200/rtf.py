"""Encoder-RTF / long-form harness for the WKV6 path, after wenet/bin/encoder-rtf.py:356-383 (the
`feats_batcher` that cuts one long recording into `chunk_size` x `batch_size` batches) and :455-550 (the timing loop and
the `.rtf` file it writes).  Same protocol, same file format, two deliberate differences:

* what runs between the two clock readings is the path this repository covers -- the attention slot of every encoder
  block, i.e. `num_blocks` bidirectional RWKV-6 time-mix blocks on the subsampled frames -- not the whole Conformer
  (feed-forward / convolution modules, CTC head and the front end are out of scope);
* every batch is bracketed by `torch.cuda.synchronize()`: the reference reads the clock without synchronising after
  `_forward_encoder` (encoder-rtf.py:497-515), which on an asynchronous device times the launches, not the work.

The input is synthetic: a recording of `minutes` minutes = 6000 feature frames per minute (10 ms hop), subsampled by 4
like the reference's Conv2dSubsampling4 (a [B, chunk/4, C] tensor per batch; the last batch is padded, its length
vector says so, exactly as feats_batcher does).

    python -m paper_accurate_fast_cheap_b200.rtf --minutes 60 --chunk-size 12000 --batch-size 8 --result-dir out/
"""
from __future__ import annotations

import argparse
import os
import time
from math import ceil
from typing import Iterator, List, Tuple

import torch

FRAMES_PER_SECOND = 100.0          # 10 ms feature hop: the reference's `num_frames / 100.0` seconds of audio
SUBSAMPLING = 4                    # Conv2dSubsampling4 (wenet/transformer/subsampling.py)


def feats_batcher(num_feats: int, chunk_size: int, batch_size: int) -> Iterator[Tuple[int, List[int]]]:
    """The reference's batching rule on frame COUNTS (encoder-rtf.py:356-383): yields (rows, lengths) per batch, lengths
    in input frames; the last batch has ceil(rest / chunk_size) rows and its last row is short by the padding."""
    per_batch = chunk_size * batch_size
    nb = ceil(num_feats / per_batch)
    for b in range(nb):
        n = min(per_batch, num_feats - b * per_batch)
        rows = batch_size
        lengths = [chunk_size] * batch_size
        if b == nb - 1:
            rows = ceil(n / chunk_size)
            lengths = [chunk_size] * rows
            lengths[-1] -= rows * chunk_size - n
        yield rows, lengths


def rtf_lines(timings: List[float], num_frames: List[int], total_elapsed: float, max_vram_mb: float, max_cpu_ram_mb: float
              ) -> List[str]:
    """The lines of the reference's `.rtf` file (encoder-rtf.py:527-550), character for character."""
    out = []
    total_compute_elapsed, total_frames = 0.0, 0
    for elapsed, nf in zip(timings, num_frames):
        local_rtf = elapsed / (nf / FRAMES_PER_SECOND)
        rate_min_per_sec = ((nf / FRAMES_PER_SECOND) / 60.0) / elapsed
        total_compute_elapsed += elapsed
        total_frames += nf
        out.append(f"num_frames={nf:8d}, elapsed={elapsed:.6f}, local_rtf={local_rtf:.6f}, {rate_min_per_sec:.2f} "
                   "minutes of audio processed per sec, compute only\n")
    total_rate_min_per_sec = ((total_frames / FRAMES_PER_SECOND) / 60.0) / total_elapsed
    out.append(f"total_frames={total_frames:8d}, total_elapsed={total_elapsed:.6f} (compute elapsed {total_compute_elapsed:.6f}), "
               f"final_rtf {total_elapsed / (total_frames / FRAMES_PER_SECOND):.6f}, {total_rate_min_per_sec:.2f} minutes of audio "
               "processed per sec\n")
    out.append(f"max_vram={max_vram_mb:.2f} MB, max_cpu_ram={max_cpu_ram_mb:.2f} MB\n")
    return out


def build_blocks(num_blocks: int, dim: int, device, bf16: bool = True):
    from . import wrappers
    blocks = []
    for layer in range(num_blocks):
        torch.manual_seed(777 + layer)
        blocks.append(wrappers.RWKV_TmixWrapper_bidirectional(64, dim, num_blocks, "v6", "bi", 2048, bf16, layer).to(device).eval())
    return blocks


def run(minutes: float, chunk_size: int, batch_size: int, num_blocks: int = 12, dim: int = 512, warmup: int = 1,
        device: str = "cuda:0", result_dir: str | None = None, name: str = "synthetic") -> List[str]:
    import psutil
    dev = torch.device(device)
    blocks = build_blocks(num_blocks, dim, dev)
    num_feats = int(round(minutes * 60 * FRAMES_PER_SECOND))
    g = torch.Generator(device=dev).manual_seed(777)
    timings, frames = [], []
    torch.cuda.reset_peak_memory_stats(dev)

    def encoder(x):
        for blk in blocks:                                   # residual connection of the encoder layer around the slot
            x = x + blk(x, x, x)[0]
        return x

    with torch.no_grad():
        t_start = None
        for i, (rows, lengths) in enumerate(feats_batcher(num_feats, chunk_size, batch_size)):
            x = torch.randn(rows, chunk_size // SUBSAMPLING, dim, device=dev, generator=g)
            if i == 0:
                for _ in range(warmup):
                    encoder(x)
                torch.cuda.synchronize(dev)
                t_start = time.time()
            torch.cuda.synchronize(dev)
            t0 = time.time()
            encoder(x)
            torch.cuda.synchronize(dev)
            timings.append(time.time() - t0)
            frames.append(int(sum(lengths)))
        total = time.time() - t_start
    lines = rtf_lines(timings, frames, total, torch.cuda.max_memory_allocated(dev) / 1024 / 1024,
                      psutil.Process().memory_info().rss / 1024 / 1024)
    if result_dir:
        os.makedirs(os.path.join(result_dir, "rtf"), exist_ok=True)
        with open(os.path.join(result_dir, "rtf", name + ".rtf"), "w") as fp:
            fp.writelines(lines)
    return lines


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--minutes", type=float, default=20.0, help="length of the synthetic recording")
    ap.add_argument("--chunk-size", type=int, default=12000, help="input frames per row (encoder-rtf.py --chunk_size)")
    ap.add_argument("--batch-size", type=int, default=8)
    ap.add_argument("--num-blocks", type=int, default=12)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--result-dir", default=None)
    a = ap.parse_args()
    print("".join(run(a.minutes, a.chunk_size, a.batch_size, a.num_blocks, a.dim, a.warmup, result_dir=a.result_dir)), end="")


if __name__ == "__main__":
    main()

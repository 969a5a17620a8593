"""Encoder-RTF harness (paper_accurate_fast_cheap_b200/rtf.py): the batching rule and the `.rtf` file format
of wenet/bin/encoder-rtf.py:356-383, 527-550."""
import re
from math import ceil

import pytest
import torch

from paper_accurate_fast_cheap_b200 import rtf


@pytest.mark.parametrize("num_feats,chunk,batch", [(360000, 12000, 8), (120000, 3000, 16), (1000, 300, 2), (299, 300, 4), (600, 300, 2)])
def test_feats_batcher_follows_the_reference_rule(num_feats, chunk, batch):
    got = list(rtf.feats_batcher(num_feats, chunk, batch))
    assert len(got) == ceil(num_feats / (chunk * batch))                       # encoder-rtf.py:357-358
    assert sum(sum(l) for _, l in got) == num_feats                            # padding is not counted as audio
    for i, (rows, lengths) in enumerate(got):
        assert rows == len(lengths) and all(0 < x <= chunk for x in lengths)
        if i < len(got) - 1:
            assert rows == batch and lengths == [chunk] * batch
        else:
            assert all(x == chunk for x in lengths[:-1])                       # only the last row of the last batch is short


def test_rtf_file_format_is_the_reference_one():
    timings, frames = [0.25, 0.125], [96000, 41000]
    lines = rtf.rtf_lines(timings, frames, 0.5, 1234.5, 678.25)
    # the reference's own f-strings (encoder-rtf.py:539, 541, 549) evaluated on the same numbers
    num_frames, elapsed = frames[0], timings[0]
    local_rtf = elapsed / (num_frames / 100.0)
    rate_min_per_sec = ((num_frames / 100.00) / 60.0) / elapsed
    assert lines[0] == f"{num_frames=:8d}, {elapsed=:.6f}, {local_rtf=:.6f}, {rate_min_per_sec:.2f} minutes of audio processed per sec, compute only\n"
    total_frames, total_elapsed, total_compute_elapsed = sum(frames), 0.5, sum(timings)
    total_rate_min_per_sec = ((total_frames / 100.00) / 60.0) / total_elapsed
    assert lines[2] == (f"{total_frames=:8d}, {total_elapsed=:.6f} (compute elapsed {total_compute_elapsed:.6f}), final_rtf "
                        f"{total_elapsed / (total_frames / 100.00):.6f}, {total_rate_min_per_sec:.2f} minutes of audio processed per sec\n")
    max_vram, max_cpu_ram = 1234.5, 678.25
    assert lines[3] == f"{max_vram=:.2f} MB, {max_cpu_ram=:.2f} MB\n"


@pytest.mark.gpu
def test_rtf_run_on_gpu(tmp_path):
    lines = rtf.run(minutes=0.5, chunk_size=1000, batch_size=2, num_blocks=2, dim=128, warmup=1, result_dir=str(tmp_path))
    assert len(lines) == 2 + 2                                                  # 3000 frames -> 2 batches, total, memory
    assert (tmp_path / "rtf" / "synthetic.rtf").read_text() == "".join(lines)
    m = re.match(r"total_frames=\s*(\d+), total_elapsed=([\d.]+) ", lines[-2])
    assert m and int(m.group(1)) == 3000 and float(m.group(2)) > 0

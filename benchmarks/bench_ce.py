"""Cross-entropy forward + backward over one LM-head chunk (4224 x 50304 bf16 logits, in place): the one-CTA-per-row
kernel (B200_CE_PIPE=0) vs the persistent bulk-copy pipeline (default), against the chunk's read + write at the measured
copy bandwidth.

    python benchmarks/bench_ce.py            # run twice, with and without B200_CE_PIPE=0
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402

rows, vocab, ld = 4224, 50257, 50304
logits = (torch.randn(rows, ld, device="cuda") * 2).to(torch.bfloat16)
src = logits.clone()
labels = torch.randint(0, vocab, (rows,), device="cuda")
loss = torch.zeros(1, device="cuda")


def run():
    k.ce_fwd_bwd(logits, labels, loss, vocab, 1.0, 1.0)


run()
torch.cuda.synchronize()
t = timeit(run)
gb = 2 * rows * ld * 2 / 1e9
print(f"B200_CE_PIPE={os.environ.get('B200_CE_PIPE', '1')}: {t * 1e3:.1f} us  ({gb / t * 1e3:.0f} GB/s of read + write)")

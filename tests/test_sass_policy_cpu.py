"""Static checks on the built library's SASS (no GPU needed: cuobjdump reads the .so).

* the hot paths are tcgen05 / TMA code (UTCHMMA, LDTM, UTMALDG, UTMASTG, UTMAREDG present; no HMMA = no mma.sync),
* the NVLS kernels carry multimem instructions,
* no TMA *load* is a multicast load: TMA multicast loads here read only this GPU's memory, since on B200 a TMA
  multicast load from another GPU's memory hung the GPU -- and this library has operands that may live in peer
  memory, so it uses none at all.  (tcgen05.commit with .multicast::cluster -- UTCBAR...MULTICAST -- is an mbarrier
  arrive, not a load, and stays.)
"""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_llm_trainer_b200", "ops", "libb200trainer.so")


@pytest.fixture(scope="module")
def sass():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe) or not os.path.exists(LIB):
        pytest.skip("cuobjdump or the built library is not available")
    return subprocess.run([exe, "-sass", LIB], capture_output=True, text=True, check=True).stdout


def _count(sass, pattern):
    return len(re.findall(pattern, sass))


def test_tensor_core_and_tma_instructions_present(sass):
    assert _count(sass, r"\bUTCHMMA") > 100 and _count(sass, r"\bUTCHMMA\.2CTA") > 50     # tcgen05.mma, CTA pairs
    assert _count(sass, r"\bLDTM") > 20 and _count(sass, r"\bSTTM") >= 1                     # tcgen05.ld / st
    assert _count(sass, r"\bUTMALDG") > 100 and _count(sass, r"\bUTMASTG") > 10              # TMA tensor loads / stores
    assert _count(sass, r"\bUTMAREDG") >= 8                                                  # TMA reduce-add (wgrad, dQ)
    assert _count(sass, r"\bUBLKCP") >= 4                                                    # 1-D bulk copies
    assert _count(sass, r"\bHMMA") == 0                                                      # no mma.sync / wmma fallback


def test_nvls_instructions_present(sass):
    assert "LDGMC" in sass or "MULTIMEM" in sass       # multimem.ld_reduce
    assert _count(sass, r"\bFFMA2") > 20                # packed fp32 math in the softmax / CE passes


def test_no_multicast_tma_loads(sass):
    loads = [ln for ln in sass.splitlines() if re.search(r"\b(UTMALDG|UBLKCP)", ln)]
    assert loads
    bad = [ln.strip() for ln in loads if re.search(r"MULTICAST|\.MC\b", ln)]
    assert not bad, bad[:5]


def test_attention_backward_uses_tma_for_its_outputs(sass):
    m = re.search(r"Function : \S*attn_bwd_kernel\S*(.*?)(?:\n\s*Function : |\Z)", sass, re.S)
    assert m, "attn_bwd_kernel not found"
    body = m.group(1)
    assert "UTMAREDG" in body and "UTMASTG" in body      # dQ reduce-add, dK / dV stores
    assert not re.search(r"\bREDG\b|\bRED\.E\.ADD", body)  # no per-thread global reductions left

"""What proves a Blackwell-native build (B200_PROFILING.md): the product library's SASS holds the 5th-generation tensor-core
and TMA instructions -- tcgen05.mma as UTC*MMA (also the CTA-pair form), tcgen05.ld/st as LDTM/STTM, cp.async.bulk as UBLKCP,
cp.async.bulk.tensor as UTMALDG -- and no legacy HMMA path.  Runs without a GPU (cuobjdump reads the cubin in the .so)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "formula-vad_b200", "libfvad_cuda.so")


@pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not on PATH")
def test_library_sass_is_tcgen05_and_tma():
    assert os.path.exists(LIB), "build the library first (__graft_entry__.build)"
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, timeout=300).stdout
    assert "sm_100a" in sass or "EF_CUDA_SM100" in sass or "arch = sm_100" in sass
    for mnemonic in ("UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UBLKCP", "UTMALDG"):
        assert mnemonic in sass, mnemonic
    # the kernels that carry them
    for kernel in ("tc_gemm_pair_kernel", "tc_gru3_kernel", "tc_gru2_kernel", "tc_gemm_kernel", "tc_gru_kernel", "stft_kernel_d1", "istft_stream_kernel",
                   "vadfft_kernel", "vadscan_warp_kernel", "eval_batch_kernel"):
        assert kernel in sass, kernel
    assert " HMMA" not in sass and "HGMMA" not in sass  # no mma.sync / wgmma fallbacks

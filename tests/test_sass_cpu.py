"""No GPU needed: the built library must contain the sm_100a machinery the design claims (tcgen05, TMA, TMEM,
cluster barriers, multimem), checked on the SASS with cuobjdump (the mnemonics are listed in
/opt/skills/guides/B200_PROFILING.md; tools/sass_report.py prints the per-kernel table)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cs744-distributed-data-parallel_b200", "_native", "libddpb200.so")


@pytest.fixture(scope="module")
def sass():
    if shutil.which("cuobjdump") is None or not os.path.exists(LIB):
        pytest.skip("cuobjdump or the built library is not available")
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, timeout=600).stdout
    kernels, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = set()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            kernels[cur].add(m.group(1))
    return kernels


def _with(kernels, name):
    return {k: v for k, v in kernels.items() if name in k}


def test_library_targets_sm100a_only():
    if shutil.which("cuobjdump") is None or not os.path.exists(LIB):
        pytest.skip("cuobjdump or the built library is not available")
    out = subprocess.run(["cuobjdump", "-lelf", LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


@pytest.mark.parametrize("kernel", ["conv3x3_tc_kernel", "wgrad_tc_kernel", "gemm_rows_tc_kernel", "conv3x3_tc2_kernel",
                                    "conv3x3_tf32_kernel", "wgrad_tf32_kernel", "first_layer_bwd_kernel"])
def test_tensor_core_kernels_use_tcgen05_tma_and_tmem(sass, kernel):
    ks = _with(sass, kernel)
    assert ks, f"no kernel named {kernel} in the library"
    for name, ops in ks.items():
        assert "UTCHMMA" in ops, (name, "no tcgen05.mma")
        assert "UTMALDG" in ops, (name, "no TMA tensor load")
        assert "LDTM" in ops, (name, "no tcgen05.ld (TMEM read)")
        assert "SYNCS" in ops, (name, "no mbarrier")
        assert not any(o.startswith("HMMA") for o in ops), (name, "legacy mma.sync in a tcgen05 kernel")


def test_wgrad_epilogue_uses_tma_reduce_add(sass):
    ks = _with(sass, "wgrad_tc_kernelILi2ELi128ELi3")
    assert ks and all("UTMAREDG" in ops for ops in ks.values())


def test_split_k_variants_use_cluster_barriers(sass):
    ks = {k: v for k, v in _with(sass, "conv3x3_tc_kernel").items() if re.search(r"Li1ELi[24]ELb[01]EEEv", k)}
    assert ks, "no cluster split-K instantiation found"
    assert all("UCGABAR_ARV" in ops for ops in ks.values())


def test_nvls_kernel_uses_multimem(sass):
    ks = _with(sass, "p2p_nvls_kernel")
    assert ks and all("LDGMC" in ops for ops in ks.values())


def test_pdl_kernels_contain_wait_and_trigger(sass):
    for kernel in ("conv3x3_tc_kernel", "bn_fwd_kernel", "bn_bwd_apply_kernel", "fc_ce_fwd_kernel"):
        ks = _with(sass, kernel)
        assert ks and all("ACQBULK" in ops and "PREEXIT" in ops for ops in ks.values()), kernel


def test_pair_tiles_issue_the_two_cta_mma(sass):
    """cta_group::2 instantiations (PAIR = true): the MMA is the 2-CTA form and the commit is multicast."""
    ks = {k: v for k, v in _with(sass, "conv3x3_tc_kernel").items() if k.endswith("ELb1EEEv14CUtensorMap_stS1_PKfPfNS_12ConvTcParamsE")}
    assert ks, "no pair instantiation found"
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, timeout=600).stdout
    for name in ks:
        body = text.split("Function : " + name, 1)[1].split("Function :", 1)[0]
        assert re.search(r"UTCHMMA\.2CTA", body), name
        assert re.search(r"UTCBAR\.2CTA\.MULTICAST", body), name


def test_block0_backward_ends_with_one_bulk_reduce_add(sass):
    """The 64 x 27 partial weight gradient leaves each CTA through cp.reduce.async.bulk (UBLKRED), the RED loop is the
    fall-back for non-contiguous destinations only."""
    ks = _with(sass, "first_layer_bwd_kernel")
    assert ks and all("UBLKRED" in ops for ops in ks.values()), {k: sorted(v & {"UBLKRED", "REDG"}) for k, v in ks.items()}

"""Build-time tier: the cross-compiled library really contains the Blackwell paths. Runs without a GPU —
nvcc builds sm_100a here and cuobjdump lists the SASS."""
import re
import shutil
import subprocess

import pytest


@pytest.fixture(scope="module")
def sass():
    if shutil.which("cuobjdump") is None or shutil.which("nvcc") is None:
        pytest.skip("CUDA toolkit not on PATH")
    from tensorflow_train_distributed_b200 import _build
    lib = _build.build_cuda()
    out = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
    per, fn = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            per[fn] = []
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Za-z0-9_.]+)", line)
        if m and fn:
            per[fn].append(m.group(1))
    assert "sm_100a" in out
    return per


def _ops(per, needle):
    hits = [ops for fn, ops in per.items() if needle in fn]
    assert hits, f"no kernel matching {needle}"
    return [o for ops in hits for o in ops]


def test_gemm_kernels_use_tcgen05_tmem_and_tma(sass):
    for k in ("gemm_group_kernel", "gemm_chain_kernel", "gemm_net_kernel"):
        ops = _ops(sass, k)
        assert any(o.startswith("UTCHMMA") for o in ops), k   # tcgen05.mma kind::f16
        assert any(o.startswith("UTMALDG") for o in ops), k   # cp.async.bulk.tensor (TMA)
        assert any(o.startswith("LDTM") for o in ops), k      # tcgen05.ld from TMEM
        assert any(o.startswith("UTCBAR") for o in ops), k    # tcgen05.commit -> mbarrier
        assert not any(o.startswith("HMMA") or o.startswith("HGMMA") for o in ops), k  # no legacy tensor path
    assert any(o == "UTMALDG.2D.MULTICAST" for o in _ops(sass, "gemm_group_kernel"))  # cluster-shared A tile
    assert any(o.startswith("ACQBULK") for o in _ops(sass, "gemm_group_kernel"))      # griddepcontrol.wait (PDL)


def test_update_kernel_has_the_nvlink_paths(sass):
    ops = _ops(sass, "param_kernelILi2E")  # kUpdate
    assert any(o.startswith("LDGMC.E.ADD.F32x4") for o in ops)   # multimem.ld_reduce.v4.f32 (NVLS, aligned groups)
    assert any(o == "LDGMC.E.ADD.F32.RN.STRONG.SYS" for o in ops)  # scalar form for unaligned groups
    assert any(o.startswith("LDG.E.128.STRONG.SYS") for o in ops)  # peer gradient loads (two-GPU path)
    assert any(o.startswith("MEMBAR.SC.SYS") for o in ops)         # release fence of the barrier-out
    push = _ops(sass, "param_kernelILi4E")  # kAsyncPush
    assert any(o.startswith("ATOMG.E.ADD.F32") and o.endswith("SYS") for o in push)  # Hogwild apply over NVLink
    assert any(o.startswith("ATOMG.E.ADD.64") and o.endswith("SYS") for o in push)   # shared global_step

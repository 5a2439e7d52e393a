"""CPU-only checks: the C-ABI library loads and exports every symbol the header declares, the host modules
refuse to run without a CUDA device, and the argument handling mirrors the reference."""
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built_lib():
    sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
    import build as bsc_build

    return bsc_build.build_lib()


def test_library_exports_every_declared_symbol(built_lib):
    import ctypes

    header = open(os.path.join(ROOT, "include", "bsc_hot.h")).read()
    declared = set(re.findall(r"^\s*(?:const\s+)?[a-z_0-9]+\*?\s+\*?(bsc_[a-z0-9_]+)\s*\(", header, re.M))
    assert {"bsc_infonce_fwd", "bsc_infonce_plan", "bsc_infonce_bwd_partials", "bsc_infonce_bwd_finish", "bsc_bank_search", "bsc_topk_merge", "bsc_l2norm_fwd"} <= declared
    lib = ctypes.CDLL(built_lib)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in bsc_hot.h but not exported"
    lib.bsc_version.restype = ctypes.c_int
    assert lib.bsc_version() == 200
    from bsc import _lib

    assert set(_lib.EXPORTS) == declared


def test_every_entry_point_is_documented_for_the_reference_maintainer():
    """INTEGRATION.md names every entry point of include/bsc_hot.h (with the reference lines it stands in for)."""
    header = open(os.path.join(ROOT, "include", "bsc_hot.h")).read()
    declared = set(re.findall(r"^\s*(?:const\s+)?[a-z_0-9]+\*?\s+\*?(bsc_[a-z0-9_]+)\s*\(", header, re.M))
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = sorted(n for n in declared if f"`{n}`" not in doc and f"`{n}(" not in doc and n not in doc)
    assert not missing, missing
    assert "loss_func.py" in header and "inference_and_eval.py" in header  # the header cites the reference


def test_sass_contains_blackwell_tensor_and_tma_ops(built_lib):
    out = subprocess.run(["cuobjdump", "-sass", built_lib], capture_output=True, text=True).stdout
    assert "UTCHMMA" in out, "tcgen05.mma missing from the SASS"
    assert "LDTM" in out, "tcgen05.ld missing from the SASS"
    assert "UTMALDG" in out, "TMA loads missing from the SASS"
    assert "HMMA.16" not in out.replace("UTCHMMA", ""), "legacy mma.sync path present"


def test_argument_errors_without_gpu(built_lib):
    import ctypes

    from bsc import _lib

    L = _lib.lib()
    rc = L.bsc_l2norm_fwd(None, 0, 4, 8, 8, None, 2, None, None, 0, None)
    assert rc == -1 and b"x is null" in L.bsc_last_error()
    nbytes = ctypes.c_size_t(0)
    assert L.bsc_bank_search_workspace_bytes(10, 100, 64, 5, 7, ctypes.byref(nbytes)) == -1
    assert L.bsc_bank_search_workspace_bytes(10, 100, 64, 5, 0, ctypes.byref(nbytes)) == 0 and nbytes.value > 0
    d = _lib.InfoNceDesc()
    d.n_global, d.n_local, d.row0, d.d, d.mode, d.world = 8, 8, 0, 128, 1, 1
    plan = _lib.InfoNcePlan()
    assert L.bsc_infonce_plan(ctypes.byref(d), _lib.BWD_AUTO, 1, 0, ctypes.byref(plan)) == -1
    assert b"Too less element" in L.bsc_last_error()
    # the plan itself needs no device: strategy and buffer sizes of BASELINE config 3 on 8 ranks
    one = ctypes.c_void_p(256)  # never dereferenced by the plan
    d.n_global, d.n_local, d.row0, d.d, d.x16_dtype, d.world, d.rank = 32768, 4096, 4096, 768, _lib.BSC_F16, 8, 1
    for m in range(3):
        d.xh[m] = one
        d.inv[m] = one
    d.r = d.rep = d.next = d.rmax = one
    assert L.bsc_infonce_plan(ctypes.byref(d), _lib.BWD_AUTO, 1, 0, ctypes.byref(plan)) == 0, L.bsc_last_error()
    assert plan.bwd_mode == _lib.BWD_SCATTER
    assert plan.exchange_bytes == 8 * 3 * 4096 * 768 * 4
    assert 3 * 2 * 4096 * 32768 <= plan.bwd_workspace_bytes <= 3 * 2 * 4096 * 32768 + (1 << 30)
    assert L.bsc_infonce_plan(ctypes.byref(d), _lib.BWD_AUTO, 0, 0, ctypes.byref(plan)) == 0
    assert plan.bwd_mode == _lib.BWD_BLOCKS and plan.exchange_bytes == 0   # no peer memory: ordered-pair blocks
    assert L.bsc_infonce_plan(ctypes.byref(d), _lib.BWD_SCATTER, 0, 0, ctypes.byref(plan)) == -2  # explicit request refused
    assert b"peer-mapped" in L.bsc_last_error()
    assert L.bsc_infonce_plan(ctypes.byref(d), _lib.BWD_AUTO, 1, 1 << 20, ctypes.byref(plan)) == 0
    assert plan.bwd_mode == _lib.BWD_RECOMPUTE                              # 1 MB budget: recompute instead of storing


def test_modules_refuse_cpu_tensors(built_lib):
    import bsc

    x = torch.randn(4, 8)
    with pytest.raises(ValueError, match="Too less element for calculating the contrastive loss."):
        bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), 1 / 0.07)(x, None, None, torch.arange(4))
    with pytest.raises(bsc.BscError, match="no CPU path"):
        bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), 1 / 0.07)(x, x, None, torch.arange(4))
    with pytest.raises(NotImplementedError):
        bsc.ClipLoss(criterion=torch.nn.CrossEntropyLoss(reduction="sum"))
    with pytest.raises(NotImplementedError):
        bsc.ClipLoss(use_horovod=True)
    if not torch.cuda.is_available():
        with pytest.raises(bsc.BscError):
            bsc.IndexFlatIP(8)


def test_resolve_mode():
    from bsc._lib import MODE_F32, MODE_TC16, MODE_TC16X3
    from bsc.loss import X3_MIN_ROWS, resolve_mode

    f32, b16 = torch.zeros(1, 256), torch.zeros(1, 256, dtype=torch.bfloat16)
    assert resolve_mode("auto", [f32, f32]) == MODE_F32
    big = torch.zeros(X3_MIN_ROWS, 256)
    assert resolve_mode("auto", [big, big]) == MODE_TC16X3          # float32 inputs, large batch: split operands
    assert resolve_mode("auto", [f32, f32], n_global=X3_MIN_ROWS) == MODE_TC16X3
    assert resolve_mode("fp32x3", [f32, f32]) == MODE_TC16X3 and resolve_mode("fp32", [big, big]) == MODE_F32
    assert resolve_mode("auto", [f32, b16]) == MODE_TC16
    assert resolve_mode("tc16", [f32]) == MODE_TC16 and resolve_mode("fp32", [b16]) == MODE_F32
    odd = torch.zeros(1, 72, dtype=torch.bfloat16)
    assert resolve_mode("auto", [odd, odd]) == MODE_F32  # widths the tensor-core kernels do not tile
    import bsc
    with pytest.raises(bsc.BscError, match="multiple of 128"):
        resolve_mode("tc16", [odd])
    with pytest.raises(ValueError):
        resolve_mode("fp8", [f32])


def test_extracted_feature_container_round_trip(tmp_path):
    """Directory container with the reference's /split/feature_type tree and labels json
    (scripts/inference_and_eval.py:804-831, :863-884)."""
    import numpy as np

    from bsc import storage

    rng = np.random.default_rng(0)

    def split(n, with_all=False):
        d = {"encoded_image_feature": rng.standard_normal((n, 8)), "encoded_dna_feature": rng.standard_normal((n, 8)),
             "encoded_language_feature": None, "averaged_feature": rng.standard_normal((n, 8)),
             "concatenated_feature": rng.standard_normal((n, 16)),
             "label_list": [{"order": f"o{i}", "family": "f", "genus": "g", "species": f"s{i}"} for i in range(n)]}
        if with_all:
            d["all_key_features"] = rng.standard_normal((3 * n, 8))
        return d

    seen, unseen, keys = split(5), split(3), split(7, True)
    root, labels = str(tmp_path / "feats"), str(tmp_path / "labels_val.json")
    storage.save_extracted_features(root, labels, seen, unseen, keys)
    s2, u2, k2 = storage.load_extracted_features(root, labels, mmap=True)
    for a, b in ((seen, s2), (unseen, u2), (keys, k2)):
        for ft in storage.ALL_TYPE_OF_FEATURES_OF_KEY:
            if a.get(ft) is None:
                assert ft not in b
            else:
                assert np.array_equal(np.asarray(b[ft]), a[ft])
        assert b["label_list"] == a["label_list"]
    assert k2["all_key_features_label"] == keys["label_list"] * 3
    import json
    assert set(json.load(open(labels))) == {"seen_gt_dict", "unseen_gt_dict", "key_gt_dict"}


def test_bench_reference_arm_prints_exactly_one_json_line():
    """bench.py contract: stdout is ONE JSON line (library chatter on fd 1 is routed to stderr); the reference arm
    runs on the host cores without a GPU."""
    import json
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout[:500]
    line = json.loads(lines[0])
    # the timed arm is the reference module itself whenever oracle/_ref was staged (build() does that here)
    from oracle import build_ref
    assert line["impl"] == "reference" and line["cpu_baseline"]["kind"] == ("reference" if build_ref.available() else "port")
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "e2e"):
        assert key in line

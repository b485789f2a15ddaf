"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol the header
declares, and refuses loudly (no CPU fallback) when there is no device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as G
    G.build()
    from transformer_stm_b200 import _lib
    return _lib


def header_symbols():
    src = open(os.path.join(ROOT, "include", "cvt_b200.h")).read()
    return sorted(set(re.findall(r"CVT_API[^;(]*?\b(cvt_\w+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound(lib):
    names = header_symbols()
    assert len(names) >= 47
    handle = C.CDLL(lib.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), "library does not export %s" % n
    assert set(names) == set(lib.exported_names()), set(names) ^ set(lib.exported_names())
    assert lib.load().cvt_abi_version() == lib.ABI_VERSION == 3


def test_default_config_is_the_reference_spec(lib):
    cfg = lib.Config()
    lib.load().cvt_default_config(C.byref(cfg))
    assert (cfg.image_h, cfg.image_w, cfg.in_chans, cfg.n_stages) == (128, 128, 1, 3)
    st = [(s.embed_dim, s.patch_size, s.stride, s.num_heads, s.qkv_method, s.with_cls_token) for s in cfg.stage[:3]]
    assert st == [(64, 7, 4, 1, 0, 0), (128, 3, 2, 2, 0, 0), (256, 3, 2, 4, 0, 1)]        # models/CvT(Par).py:66-72
    assert (cfg.proc_dim, cfg.proc_hidden, cfg.num_classes, cfg.mlp_ratio) == (5, 256, 1, 4)
    assert abs(cfg.drop_rate - 0.1) < 1e-7 and abs(cfg.ln_eps - 1e-6) < 1e-12 and abs(cfg.bn_eps - 1e-3) < 1e-9


def test_python_config_struct_matches_c_layout(lib):
    import transformer_stm_b200 as T
    c = T.CvtConfig.reference().to_c()
    d = lib.Config()
    lib.load().cvt_default_config(C.byref(d))
    c.max_batch = d.max_batch
    assert bytes(c) == bytes(d)


def test_no_cpu_fallback_without_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import transformer_stm_b200 as T
    with pytest.raises(T.CvtError) as e:
        T.CvtEngine(T.CvtConfig.reference())
    assert "no CUDA device" in str(e.value) or "CPU fallback" in str(e.value)


def test_invalid_configs_rejected_before_touching_the_device(lib):
    import transformer_stm_b200 as T
    with pytest.raises(T.CvtError) as e:
        T.CvtEngine(T.CvtConfig.reference(dims=(64, 100, 256)))
    assert "embed_dim" in str(e.value)
    with pytest.raises(T.CvtError):
        T.CvtEngine(T.CvtConfig.reference(max_batch=0))


def test_product_does_not_import_the_oracle():
    prod = os.path.join(ROOT, "transformer-stm_b200")
    pat = re.compile(r"^\s*(from\s+oracle|import\s+oracle|from\s+\.\.?oracle)", re.M)
    for dp, _, files in os.walk(prod):
        for f in files:
            if f.endswith(".py"):
                assert not pat.search(open(os.path.join(dp, f)).read()), (dp, f)

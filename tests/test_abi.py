"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/yupvq.h declares, and the config mapping matches the reference (no GPU needed)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "yupvq.h")).read()
    return sorted(set(re.findall(r"YUPVQ_API\s+[\w\s\*]+?\b(yupvq_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from yuptexture_b200 import binding
    assert os.path.exists(binding.LIB_PATH), "libyupvq.so not built (python -c 'import __graft_entry__ as g; g.build()')"
    lib = ctypes.CDLL(binding.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/yupvq.h but not exported: {missing}"


def test_codebook_size_matches_reference_mapping(golden_colour):
    from yuptexture_b200 import binding
    got = [binding.codebook_size(float(q)) for q in golden_colour["codebook_size_q"]]
    assert got == golden_colour["codebook_size"].tolist()
    assert binding.codebook_size(0.8) == 512


def test_config_init_mirrors_compress():
    from yuptexture_b200 import binding
    cfg = binding.make_config(quality=0.8, max_iterations=32, seed=42)
    assert (cfg.codebook_size, cfg.max_iterations, cfg.metric, cfg.seed) == (512, 32, binding.METRIC_OKLAB, 42)
    assert cfg.sample_ratio == 1.0


def test_struct_layouts_match_header():
    from yuptexture_b200 import binding
    assert ctypes.sizeof(binding.Config) == 32
    assert ctypes.sizeof(binding.Stats) == 64


def test_no_device_fails_loudly():
    """Without a CUDA device the product must refuse to run (no CPU fallback)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from yuptexture_b200 import binding
    with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
        binding.Encoder(0)

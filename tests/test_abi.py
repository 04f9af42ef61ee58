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


def test_headers_compile_as_plain_c(tmp_path):
    """The boundary is a C ABI: include/yupvq.h and include/yupcodec.h must compile as C99 (extern "C" only under C++)."""
    import subprocess
    src = tmp_path / "abi.c"
    src.write_text('#include "yupvq.h"\n#include "yupcodec.h"\nint main(void) { yupvq_config c; (void)c; return 0; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_cpp_dropin_headers_compile_standalone(tmp_path):
    """The reference-named headers (vq_bcn_compressor.h, vq_encoder.h, bcn_compressor.h, vq_bcn_types.h) compile on their own,
    with the reference's class and field names."""
    import subprocess
    for h in ("vq_bcn_compressor.h", "vq_encoder.h", "bcn_compressor.h", "vq_bcn_types.h"):
        src = tmp_path / (h + ".cpp")
        src.write_text('#include "%s"\nint main() { CompressionParams p; p.vq_maxIterations = 1; VQCodebook cb; (void)cb; return (int)p.alphaThreshold - 128; }\n' % h)
        r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
        assert r.returncode == 0, (h, r.stderr)


def test_host_library_exports_its_c_entry_points():
    from yuptexture_b200 import hostapi
    assert hostapi.available(), "libyuptexture.so not built"
    lib = ctypes.CDLL(hostapi.LIB_PATH)
    for name in ("yuptexture_codec_name", "yuptexture_load_codec", "yuptexture_compress_timed", "yuptexture_decompress_to_bcn",
                 "yuptexture_set_gpu_bcn_encode", "yuptexture_last_encoder_name"):
        assert hasattr(lib, name), name


def test_product_never_touches_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import or link oracle/ (tier rule 3)."""
    pkg = os.path.join(ROOT, "yuptexture_b200")
    offenders = []
    for base, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                continue
            text = open(os.path.join(base, f), errors="ignore").read()
            if re.search(r"^\s*(from|import)\s+oracle\b|orcbind|refbind|libvqoracle|libyupref", text, re.M):
                offenders.append(os.path.relpath(os.path.join(base, f), ROOT))
    assert not offenders, offenders

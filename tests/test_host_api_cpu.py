"""CPU-side checks of the drop-in C++ API: headers compile stand-alone with the reference's names, struct
layouts match the reference's stream format, and libyuptexture.so exports the classes."""
import os
import subprocess

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def test_reference_named_headers_compile(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text('''
#include "vq_bcn_types.h"
#include "bcn_compressor.h"
#include "vq_encoder.h"
#include "vq_bcn_compressor.h"
#include <cstddef>
static_assert(sizeof(TextureInfo) == 24, "header size");
static_assert(offsetof(TextureInfo, format) == 8 && offsetof(TextureInfo, originalChannelCount) == 12, "offsets");
static_assert(offsetof(TextureInfo, storedCodebookEntries) == 16 && offsetof(TextureInfo, compressionFlags) == 20, "offsets");
static_assert(sizeof(CompressionParams) == 44, "CompressionParams layout (reference include/vq_bcn_types.h:110-126)");
int main() { CompressionParams p; return p.vq_maxIterations == 32 && p.numThreads == 16 && p.bcFormat == BCFormat::BC7 ? 0 : 1; }
''')
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)])


def test_host_library_exports_classes():
    lib = os.path.join(ROOT, "yuptexture_b200", "libyuptexture.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "yuptexture_b200", "host")])
    syms = subprocess.run(["nm", "-DC", lib], capture_output=True, text=True).stdout
    for want in ("VQBCnCompressor::Compress(", "VQBCnCompressor::CompressHDR(", "VQBCnCompressor::DecompressToBCn(",
                 "VQEncoder::BuildCodebook(", "VQEncoder::QuantizeBlocks(", "VQEncoder::BuildCodebookHDR(",
                 "VQEncoder::QuantizeBlocksHDR(", "CompressionConfig::SetQuality(", "BCnCompressor::Decompress("):
        assert want in syms, want

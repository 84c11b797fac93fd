"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/leadyolo.h declares,
argument validation works without a GPU, the packed-weight layout agrees between C and Python, and the tcgen05
descriptor encoders agree with CUTLASS' cute::UMMA when its headers are available."""
import ctypes as C
import glob
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "leadyolo.h")).read()
    return sorted(set(re.findall(r"\b(ly_[a-z0-9_]+)\s*\(", hdr)))


def test_every_declared_symbol_is_exported():
    from lead_yolo_b200 import _lib
    lib = _lib.lib()
    names = _declared()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.ly_version() >= 100


def test_python_signature_table_matches_header():
    from lead_yolo_b200 import lowlevel as ll
    hdr = open(os.path.join(ROOT, "include", "leadyolo.h")).read()
    for name, sig in ll._SIGS.items():
        m = re.search(r"int " + name + r"\((.*?)\);", hdr, re.S)
        assert m, name
        exp = ""
        for p in [q.strip() for q in m.group(1).replace("\n", " ").split(",")]:
            if "ly_stream_t" in p:
                continue
            exp += "p" if "*" in p else "l" if p.startswith("long") else "f" if p.startswith("float") else "i"
        assert exp == sig, (name, sig, exp)


def test_conv_descriptor_validation_and_chunk_count():
    from lead_yolo_b200 import _lib
    from lead_yolo_b200.packing import pack_conv_weight, packed_k, unpack_conv_weight
    lib = _lib.lib()
    d = _lib.ConvDesc()
    d.n_src = 2
    d.src[0] = d.src[1] = 1024
    d.src_c[0], d.src_c[1] = 48, 24
    d.src_ld[0], d.src_ld[1] = 48, 24
    d.N, d.H, d.W = 2, 20, 20
    d.kh = d.kw = 3
    d.stride, d.pad, d.Cout = 1, 1, 24
    assert lib.ly_conv_num_chunks(C.byref(d)) == packed_k([48, 24], 3, 3, False) // 64 == 18
    d.kh = d.kw = 2
    d.stride, d.pad = 2, 0
    d.n_src = 1
    d.src_c[0] = d.src_ld[0] = 24
    assert lib.ly_conv_num_chunks(C.byref(d)) == packed_k([24], 2, 2, True) // 64 == 2
    d.kh = d.kw = 5
    assert lib.ly_conv_num_chunks(C.byref(d)) == -2          # LY_EUNSUPPORTED
    assert b"not supported" in lib.ly_last_error()
    d.kh = d.kw = 1
    d.stride = 1
    d.src_c[0] = 12                                           # not a multiple of 8
    assert lib.ly_conv_num_chunks(C.byref(d)) == -1          # LY_EINVAL
    # pack / unpack round trip (weight-gradient path)
    for (co, srcs, k, patch) in ((24, [48, 24], 3, False), (40, [24], 2, True), (18, [128], 1, False)):
        w = torch.randn(co, sum(srcs), k, k)
        wp = pack_conv_weight(w, srcs, patch=patch)
        assert wp.shape == (co, packed_k(srcs, k, k, patch)) and wp.dtype == torch.bfloat16
        back = unpack_conv_weight(wp.float(), co, srcs, k, k, patch)
        assert torch.equal(back, w.bfloat16().float())


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from lead_yolo_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(_lib.LeadYoloError, match="no CPU / PyTorch fallback"):
        _lib.lib()


def test_modules_refuse_cpu_tensors():
    """No CPU fallback: a forward on CPU tensors must raise, not silently compute something."""
    from lead_yolo_b200 import modules as M
    m = M.Conv(16, 16, 1, 1).eval()
    with torch.no_grad(), pytest.raises(Exception):
        m(torch.zeros(1, 16, 8, 8))


def _cutlass_include():
    for pat in ("flashinfer/data/cutlass/include", "tilelang/3rdparty/cutlass/include"):
        for sp in sys.path:
            p = os.path.join(sp, pat)
            if os.path.isdir(p):
                return p
    hits = glob.glob("/opt/*/.venv/lib/python3*/site-packages/flashinfer/data/cutlass/include")
    return hits[0] if hits else None


@pytest.mark.skipif(_cutlass_include() is None, reason="CUTLASS headers not available")
def test_umma_descriptor_encoders_match_cutlass(tmp_path):
    src = tmp_path / "chk.cu"
    src.write_text(r'''
#include <cstdio>
#include <cute/arch/mma_sm100_desc.hpp>
#include <cute/numeric/numeric_types.hpp>
#include "%s/lead_yolo_b200/csrc/ly_tc.cuh"
int main() {
  using namespace cute;
  int bad = 0;
  bad |= (uint32_t)UMMA::make_instr_desc<bfloat16_t, bfloat16_t, float, 128, 256, UMMA::Major::K, UMMA::Major::K>() !=
         ly::umma_idesc_bf16_f32(128, 256, 0, 0);
  bad |= (uint32_t)UMMA::make_instr_desc<bfloat16_t, bfloat16_t, float, 128, 48, UMMA::Major::K, UMMA::Major::K>() !=
         ly::umma_idesc_bf16_f32(128, 48, 0, 0);
  bad |= (uint32_t)UMMA::make_instr_desc<bfloat16_t, bfloat16_t, float, 128, 192, UMMA::Major::MN, UMMA::Major::MN>() !=
         ly::umma_idesc_bf16_f32(128, 192, 1, 1);
  UMMA::SmemDescriptor s; s.desc_ = 0; s.version_ = 1; s.lbo_mode_ = 0;
  s.layout_type_ = uint8_t(UMMA::LayoutType::SWIZZLE_128B); s.start_address_ = (0x12340 >> 4); s.base_offset_ = 0;
  s.stride_byte_offset_ = 64; s.leading_byte_offset_ = 1;
  bad |= s.desc_ != ly::umma_smem_desc_sw128(0x12340, 16, 1024);
  s.leading_byte_offset_ = 1024;
  bad |= s.desc_ != ly::umma_smem_desc_sw128(0x12340, 16384, 1024);
  printf(bad ? "MISMATCH\n" : "MATCH\n");
  return bad;
}
''' % ROOT)
    exe = tmp_path / "chk"
    r = subprocess.run(["nvcc", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I", _cutlass_include(), "-o", str(exe),
                        str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and "MATCH" in r.stdout


def test_ctypes_struct_mirrors_match_the_header(tmp_path):
    """ly_conv_desc / ly_loss_desc / ly_pack_job as gcc lays them out from include/leadyolo.h against the ctypes mirrors in
    lead_yolo_b200/_lib.py: same size, same offset for every field (an ABI drift here corrupts launches silently)."""
    from lead_yolo_b200 import _lib
    pairs = [("ly_conv_desc", _lib.ConvDesc), ("ly_loss_desc", _lib.LossDesc), ("ly_pack_job", _lib.PackJob)]
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "leadyolo.h"', "int main(void) {"]
    for cname, cls in pairs:
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    got = {}
    for ln in out.strip().splitlines():
        s, f, v = ln.split()
        got[(s, f)] = int(v)
    for cname, cls in pairs:
        assert got[(cname, "size")] == C.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert got[(cname, fname)] == getattr(cls, fname).offset, (cname, fname)

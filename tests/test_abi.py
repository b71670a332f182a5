"""CPU tests of the drop-in boundary: libmtg_gpu.so loads without a GPU and exports every
symbol include/mtg_gpu.h declares; the ctypes structs match the C header; entry points
fail loudly (no CPU fallback) when there is no CUDA device. No compute calls here."""
import ctypes
import os
import re
import subprocess

import pytest

import mtg_b200
from mtg_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mtg_gpu.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mtg_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_reference_interface():
    names = declared_functions()
    # IBOSSChunkConstructor::initialize / add_sequences / build_chunk (boss_chunk_construct.hpp:18-34)
    for must in ("mtg_create", "mtg_destroy", "mtg_push_reads", "mtg_build_chunk", "mtg_free_chunk",
                 "mtg_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    for name in declared_functions():
        assert hasattr(lib, name), "libmtg_gpu.so does not export " + name
    assert sorted(_lib.EXPORTS) == sorted(n for n in declared_functions())


def test_struct_layout_matches_header(tmp_path):
    """sizeof/offsetof of the C structs == the ctypes mirrors (compiled with gcc as plain C,
    which also proves the header is valid C, not only C++)."""
    fields_chunk = [f[0] for f in _lib.MtgChunk._fields_]
    fields_prm = [f[0] for f in _lib.MtgParams._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "mtg_gpu.h"', 'int main(void){',
            'printf("%zu\\n", sizeof(mtg_params));', 'printf("%zu\\n", sizeof(mtg_chunk));']
    for f in fields_prm:
        prog.append('printf("%%zu\\n", offsetof(mtg_params, %s));' % f)
    for f in fields_chunk:
        prog.append('printf("%%zu\\n", offsetof(mtg_chunk, %s));' % f)
    prog.append("return 0;}")
    c = tmp_path / "layout.c"
    c.write_text("\n".join(prog))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert out[0] == ctypes.sizeof(_lib.MtgParams)
    assert out[1] == ctypes.sizeof(_lib.MtgChunk)
    off = out[2:]
    for i, f in enumerate(fields_prm):
        assert off[i] == getattr(_lib.MtgParams, f).offset, f
    for i, f in enumerate(fields_chunk):
        assert off[len(fields_prm) + i] == getattr(_lib.MtgChunk, f).offset, f


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device behaviour")
def test_no_cpu_fallback_without_a_device():
    with pytest.raises(mtg_b200.MtgError) as ei:
        mtg_b200.BOSSChunkConstructor.initialize(30, both_strands=True)
    assert ei.value.code == 2   # MTG_ERR_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_argument_validation_precedes_device_probing():
    # boss_chunk_construct.cpp:1057-1062 / :1157-1159: the same rejections as the reference
    for kw in (dict(k=0), dict(k=85), dict(k=20, bits_per_count=33),
               dict(k=20, both_strands=True, alphabet="Protein"), dict(k=51, alphabet="Protein")):
        k = kw.pop("k")
        with pytest.raises(mtg_b200.MtgError) as ei:
            mtg_b200.BOSSChunkConstructor.initialize(k, **kw)
        assert ei.value.code == 1, kw   # MTG_ERR_BAD_ARG

"""CPU tests of the drop-in boundary: libtnt.so loads and exports every symbol include/tnt.h declares,
struct layouts agree with the header, and the product refuses to run without a GPU (no fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from tests.util import pkg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = pkg.model._lib


def _header_functions():
    src = open(os.path.join(ROOT, "include", "tnt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tnt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _header_functions()
    assert len(names) >= 20
    lib = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"libtnt.so does not export {n}"
    assert sorted(_lib.SIGNATURES) == names, "ctypes signatures and include/tnt.h disagree"


def test_struct_layouts_match_header():
    # compile a tiny C program against the header and compare sizeof/offsetof with the ctypes mirrors
    prog = r'''
    #include <stdio.h>
    #include <stddef.h>
    #include "tnt.h"
    int main(void) {
      printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(tnt_config), offsetof(tnt_config, dropout_input),
             offsetof(tnt_config, clipnorm), sizeof(tnt_var_info), offsetof(tnt_var_info, offset),
             sizeof(tnt_metrics), offsetof(tnt_metrics, l2), sizeof(tnt_dense_config),
             offsetof(tnt_dense_config, clipnorm), offsetof(tnt_config, attn_loss), offsetof(tnt_config, agc_eps),
             offsetof(tnt_metrics, attn_elem_mse));
      return 0; }'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "p.c"), "w").write(prog)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "p.c"), "-o",
                               os.path.join(d, "p")])
        got = [int(x) for x in subprocess.check_output([os.path.join(d, "p")]).split()]
    want = [C.sizeof(_lib.TntConfig), _lib.TntConfig.dropout_input.offset, _lib.TntConfig.clipnorm.offset,
            C.sizeof(_lib.TntVarInfo), _lib.TntVarInfo.offset.offset, C.sizeof(_lib.TntMetrics),
            _lib.TntMetrics.l2.offset, C.sizeof(_lib.TntDenseConfig), _lib.TntDenseConfig.clipnorm.offset,
            _lib.TntConfig.attn_loss.offset, _lib.TntConfig.agc_eps.offset, _lib.TntMetrics.attn_elem_mse.offset]
    assert got == want


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    groups = pkg.synthetic.small_groups(200, 6, seed=1)
    m = pkg.NIC((groups, [8] * 6), 16, 16, 16, 8, 37, 5, 0, 0, 0, 0, 0, 0, 0.01, 0.001, 3e-5, 1e-5)
    with pytest.raises((_lib.TntError, RuntimeError, AssertionError, ValueError)):
        m.build(200, 4)


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(ImportError):
        _lib.load(str(tmp_path / "libtnt.so"))


def test_synthetic_contract():
    syn = pkg.synthetic
    g, out = syn.select_groups(32)
    assert len(g) == 345 and out == [32] * 345
    assert all(x.dtype == np.int32 and x.min() >= 0 and x.max() < syn.N_VERTICES for x in g)
    assert all((np.diff(x) > 0).all() for x in g)             # np.where-style sorted indices
    (x, cap, a0, c0), tgt = syn.make_batch(8, seed=3, n_vertices=1000, onehot=True)
    assert x.shape == (8, 1000) and cap.shape == (8, 15) and a0.shape == (8, 512) and tgt.shape == (8, 15, 5001)
    assert (cap[:, 0] == syn.START_ID).all() and (tgt.argmax(-1)[:, :-1] == cap[:, 1:]).all()
    assert (tgt.argmax(-1)[:, -1] == 0).all()
    # the literal reference indexing overlaps hemispheres (load_avg_betas.py:64-72)
    g2 = syn.make_groups(hemi_offset=False)
    assert max(x.max() for x in g2) < syn.N_VERTICES // 2


def test_image_groups_are_contiguous_rows():
    g = pkg.image_groups(4, 5)
    assert [list(x) for x in g][1] == [5, 6, 7, 8, 9]


def test_batch_assembly_entry_points_validate_arguments_before_touching_cuda():
    # argument checks come first, so this runs without a GPU: negative status + message, nothing aborts
    lib = _lib.load()
    assert lib.tnt_assemble_betas(None, 0, 0, None, 0, None, None, None) == -1
    assert b"tnt_assemble_betas" in lib.tnt_data_last_error()
    assert lib.tnt_assemble_captions(None, 0, 0, None, 0, None, None, None, 0, None, None) == -1
    assert b"tnt_assemble_captions" in lib.tnt_data_last_error()
    assert lib.tnt_data_launch_count() == 0

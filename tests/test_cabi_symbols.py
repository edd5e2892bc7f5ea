"""CPU-side checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and
exports every symbol include/tha_b200.h declares (no compute calls here: no GPU)."""
import os

from pkgload import load_pkg


def test_library_exports_every_declared_symbol():
    build = load_pkg("build")
    so = build.build_cabi()
    assert os.path.exists(so)
    cabi = load_pkg("cabi")
    names = cabi.declared_symbols()
    assert len(names) >= 40 and "tha_forward_batch_f32" in names and "tha_decode_step_q8" in names
    lib = cabi.Lib(so)
    assert lib.missing_symbols() == []
    assert lib.c.tha_version() >= 100


def test_struct_layouts_match_the_reference_structs():
    import ctypes as C
    cabi = load_pkg("cabi")
    assert C.sizeof(cabi.Config) == 28                      # 7 ints, fread from the file (src/utils.cpp:155)
    assert C.sizeof(cabi.WeightsF32) == 12 * C.sizeof(C.c_void_p)
    assert C.sizeof(cabi.State) == 16 * C.sizeof(C.c_void_p)
    assert [f[0] for f in cabi.WeightsF32._fields_][:4] == ["token_embedding_table", "rms_att_weight", "rms_ffn_weight", "wq"]


def test_sass_is_sm100a_only():
    import subprocess
    build = load_pkg("build")
    so = build.build_cabi()
    out = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    archs = {l.split(".")[-2] for l in out.splitlines() if ".cubin" in l}
    assert archs == {"sm_100a"}, archs

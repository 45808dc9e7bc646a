"""CPU-side checks of the drop-in boundary: the header, the ctypes table and both shared
libraries agree on the exported entry points (no compute calls -- there is no GPU here)."""
import ctypes
import pathlib
import re
import subprocess

import pytest

from opal_b200 import abi

ROOT = pathlib.Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "opal_b200.h").read_text()


def declared():
    return sorted(set(re.findall(r"OPAL_FN\((\w+)\)\s*\(", HEADER)) - {"name"})


def test_header_and_binding_table_agree():
    assert declared() == sorted(abi.SIGNATURES)


def test_header_compiles_as_c99(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "opal_b200.h"\nint main(void){ return sizeof(opal_migrant) == 128 ? 0 : 1; }\n')
    exe = tmp_path / "t"
    subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 0


def test_product_library_exports_every_symbol():
    so = ROOT / "opal_b200" / "csrc" / "libopal_b200.so"
    if not so.exists():
        subprocess.check_call(["make", "-C", str(so.parent)], stdout=subprocess.DEVNULL)
    lib = ctypes.CDLL(str(so))
    for name in declared():
        assert hasattr(lib, "opal_" + name), name
    # binding the whole table must succeed too (argtypes/restype set on every entry point)
    abi.OpalABI(lib, "opal_")


def test_product_has_no_cpu_fallback():
    """Without a CUDA device opal_create must fail with OPAL_ECUDA, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from opal_b200 import lib
    a = lib.load()
    h = ctypes.c_void_p()
    assert a.create(ctypes.byref(h), 0, 64, 0.0, 1.0, 1e-9, 0, 0, 0, 1, 0, 0) == abi.ECUDA
    with pytest.raises(abi.OpalError):
        abi.Slab(a, 64, 0.0, 1.0)


def test_product_never_references_the_oracle():
    for path in list((ROOT / "opal_b200").rglob("*.py")) + list((ROOT / "opal_b200" / "csrc").glob("*.cu")) + list((ROOT / "opal_b200" / "csrc").glob("*.h")):
        text = path.read_text()
        assert "opal_ref_" not in text and "libopal_oracle" not in text, path


def test_oracle_exports_every_symbol(oracle_cdll):
    for name in declared():
        assert hasattr(oracle_cdll, "opal_ref_" + name), name


def test_integration_md_declares_every_symbol():
    """INTEGRATION.md's extern "C" block (the reference-side binding a maintainer adds) lists every entry point"""
    text = (ROOT / "INTEGRATION.md").read_text()
    bound = set(re.findall(r"pub fn opal_(\w+)\(", text))
    assert sorted(bound) == declared()


def test_plain_c_caller(tmp_path, oracle_cdll):
    """examples/call_from_c.c: the ABI used from C99 with nothing but the header.  Over the oracle it runs 50 steps and keeps
    its particles and energy; linked against the product library on a box without a GPU it stops at opal_create
    (OPAL_ECUDA) -- the same source, no fallback."""
    src = ROOT / "examples" / "call_from_c.c"
    common = ["/usr/bin/gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", f"-I{ROOT / 'include'}", str(src)]
    ref = tmp_path / "call_ref"
    subprocess.check_call(common + ["-DOPAL_PREFIX=opal_ref_", f"-L{ROOT / 'oracle'}", "-lopal_oracle", f"-Wl,-rpath,{ROOT / 'oracle'}", "-lm", "-o", str(ref)])
    out = subprocess.run([str(ref)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "particles 16384" in out.stdout and "total energy drift" in out.stdout
    so = ROOT / "opal_b200" / "csrc"
    gpu = tmp_path / "call_gpu"
    subprocess.check_call(common + [f"-L{so}", "-lopal_b200", f"-Wl,-rpath,{so}", "-lm", "-o", str(gpu)])
    import torch
    if not torch.cuda.is_available():
        out = subprocess.run([str(gpu)], capture_output=True, text=True, timeout=120)
        assert out.returncode == 1 and "status 2" in out.stderr


def test_integration_md_arities_match_the_binding_table():
    """every `pub fn opal_*` of INTEGRATION.md takes as many arguments as the ctypes table (and hence the header) says,
    with pointers where the table has pointers"""
    text = (ROOT / "INTEGRATION.md").read_text()
    import ctypes as C
    for name, args in re.findall(r"pub fn opal_(\w+)\(([^)]*)\)", text):
        params = [a.strip() for a in args.split(",") if a.strip()]
        want = abi.SIGNATURES[name]
        assert len(params) == len(want), (name, params, len(want))
        for p, t in zip(params, want):
            is_ptr_rust = "*" in p.split(":", 1)[1]
            is_ptr_c = t is C.c_void_p or t is C.c_char_p or hasattr(t, "contents") or getattr(t, "_type_", None) == "P"
            assert is_ptr_rust == is_ptr_c, (name, p, t)


def test_header_arities_match_the_binding_table():
    """the C prototypes of include/opal_b200.h and the ctypes table agree argument by argument on count and pointer-ness"""
    import ctypes as C
    text = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    protos = re.findall(r"OPAL_FN\((\w+)\)\s*\(([^;]*?)\)\s*;", text, flags=re.S)
    assert sorted(n for n, _ in protos) == declared()
    for name, args in protos:
        params = [a.strip() for a in args.replace("\n", " ").split(",") if a.strip() and a.strip() != "void"]
        want = abi.SIGNATURES[name]
        assert len(params) == len(want), (name, params, len(want))
        for p, t in zip(params, want):
            is_ptr_h = "*" in p or "[" in p
            is_ptr_c = t is C.c_void_p or t is C.c_char_p or hasattr(t, "contents") or getattr(t, "_type_", None) == "P"
            assert is_ptr_h == is_ptr_c, (name, p, t)

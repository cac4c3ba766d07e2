"""The C header, the ctypes binding table and the built library must describe the same ABI.

`include/b200rnn.h` is what a C / C++ host compiles against, `pytorch_distributed_rnn_b200/_native.py` restates every
prototype and struct for ctypes, and nothing but convention keeps the two in step: a field added on one side only would
shift every later field of a struct that is passed by pointer to the kernels' host code (the communicator context, the
model descriptor).  Checked here, on the CPU:

* every function the binding declares has a prototype in the header with the same number of parameters, the same
  parameter classes (pointer / int32 / int64 / uint32 / float) and the same result class, and is exported by the library;
* every prototype of the header is exported by the library (nothing declared but never defined);
* the ctypes structures have the sizes and field offsets a C compiler gives the header's structs.
"""
import ctypes as C
import re
import shutil
import subprocess
from pathlib import Path

import pytest

from pytorch_distributed_rnn_b200 import _native

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "b200rnn.h"


def _strip_comments(text):
    return re.sub(r"/\*.*?\*/", " ", text, flags=re.S)


def _prototypes():
    """{name: (result class, [parameter classes])} parsed from the header."""
    text = _strip_comments(HEADER.read_text())
    protos = {}
    for m in re.finditer(r"(?m)^\s*((?:const\s+)?[A-Za-z_][A-Za-z0-9_]*\s*\**)\s*(b200rnn_\w+)\s*\(([^;{}]*?)\)\s*;", text):
        res, name, params = m.group(1), m.group(2), m.group(3)
        params = [p.strip() for p in params.replace("\n", " ").split(",")]
        if params == ["void"] or params == [""]:
            params = []
        protos[name] = (_classify(res), [_classify(p) for p in params])
    return protos


def _classify(decl):
    decl = decl.strip()
    if "*" in decl:
        return "ptr"
    base = decl.replace("const", "").split()
    t = base[0] if base else ""
    return {"int32_t": "i32", "int": "i32", "int64_t": "i64", "uint32_t": "u32", "float": "f32", "void": "void"}.get(t, t)


def _ctype_class(t):
    if t is None:
        return "void"
    if t in (C.c_void_p, C.c_char_p) or hasattr(t, "contents") or (isinstance(t, type) and issubclass(t, C._Pointer)):
        return "ptr"
    return {C.c_int32: "i32", C.c_int64: "i64", C.c_uint32: "u32", C.c_float: "f32", C.c_uint64: "u64"}.get(t, str(t))


@pytest.fixture(scope="module")
def lib():
    if not _native.available():
        pytest.skip("libb200rnn.so is not built and nvcc is not available")
    return _native.lib()


def test_binding_table_matches_header_prototypes(lib):
    protos = _prototypes()
    assert len(protos) > 80, "header parse lost most prototypes"
    bound = {n: getattr(lib, n) for n in dir(lib) if n.startswith("b200rnn_")}
    bound = {n: f for n, f in bound.items() if f.argtypes is not None}
    assert len(bound) > 70
    for name, fn in sorted(bound.items()):
        assert name in protos, f"{name} is bound in _native.py but has no prototype in include/b200rnn.h"
        res, params = protos[name]
        got = [_ctype_class(t) for t in fn.argtypes]
        assert got == params, f"{name}: header says {params}, _native.py says {got}"
        assert _ctype_class(fn.restype) == res, f"{name}: result {res} in the header, {_ctype_class(fn.restype)} in _native.py"


def test_every_prototype_is_exported(lib):
    missing = [n for n in _prototypes() if not hasattr(lib, n)]
    assert not missing, f"declared in include/b200rnn.h but not defined in libb200rnn.so: {missing}"


STRUCTS = {
    "b200rnn_model_desc": _native.ModelDesc,
    "b200rnn_optim_desc": _native.OptimDesc,
    "b200rnn_comm_ctx": _native.CommCtx,
    "b200rnn_topology": _native.Topology,
}


def test_struct_layouts_match_a_c_compiler(tmp_path):
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void) {"]
    for cname, ct in STRUCTS.items():
        lines.append(f'  printf("{cname} sizeof %zu\\n", sizeof({cname}));')
        for fname, *_ in ct._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run([cc, "-std=c99", "-Wall", "-Werror", str(src), "-o", str(exe)], check=True)   # the header is plain C
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    want = {}
    for line in out.splitlines():
        cname, field, value = line.split()
        want[(cname, field)] = int(value)
    for cname, ct in STRUCTS.items():
        assert C.sizeof(ct) == want[(cname, "sizeof")], f"sizeof({cname}): C {want[(cname, 'sizeof')]}, ctypes {C.sizeof(ct)}"
        for fname, *_ in ct._fields_:
            assert getattr(ct, fname).offset == want[(cname, fname)], f"{cname}.{fname}: C offset {want[(cname, fname)]}, ctypes {getattr(ct, fname).offset}"

"""CPU: the C-ABI library loads and exports every symbol include/nkbs.h declares (no compute calls)."""
import ctypes
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    text = (ROOT / "include" / "nkbs.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nkbs_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    names = _declared()
    for n in ("nkbs_version", "nkbs_last_error", "nkbs_encode_fwd", "nkbs_encode_bwd", "nkbs_mlp_fwd", "nkbs_mlp_bwd",
              "nkbs_pdf_sample", "nkbs_composite_fwd", "nkbs_composite_bwd", "nkbs_adam_step"):
        assert n in names


def test_library_exports_every_declared_symbol():
    from nerf_kbs_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(str(ROOT / "nerf_kbs_b200" / "libnkbs.so"))
    for n in _declared():
        assert hasattr(lib, n), f"libnkbs.so does not export {n}"
    lib.nkbs_version.restype = ctypes.c_int
    assert lib.nkbs_version() >= 100
    # the ctypes table binds exactly the declared compute entry points
    assert set(_lib.SIGNATURES) == set(_declared()) - {"nkbs_version", "nkbs_last_error"}


def test_cpu_tensors_are_refused():
    """No CPU fallback: operators raise on CPU tensors instead of computing something else."""
    import pytest
    import torch
    import nerf_kbs_b200 as nk
    enc = nk.HashEncoding(num_levels=2, max_res=32, log2_hashmap_size=8)
    with pytest.raises(RuntimeError):
        enc(torch.rand(4, 3))


def _prototypes():
    """name -> list of parameter declarations, parsed from include/nkbs.h."""
    text = (ROOT / "include" / "nkbs.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(?:int|const char\s*\*)\s+(nkbs_[a-z0-9_]+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        params = [p.strip() for p in m.group(2).replace("\n", " ").split(",")]
        out[m.group(1)] = [] if params == ["void"] else params
    return out


def test_ctypes_signatures_match_the_header_prototypes():
    """Every binding in _lib.SIGNATURES has as many arguments as the C prototype, pointer arguments bound as pointers,
    64-bit sizes as int64 and scalars as their C type: an ABI change that forgets one side fails here, on the CPU."""
    from nerf_kbs_b200 import _lib
    protos = _prototypes()
    assert set(protos) - {"nkbs_version", "nkbs_last_error"} == set(_lib.SIGNATURES)
    for name, argtypes in _lib.SIGNATURES.items():
        params = protos[name]
        assert len(params) == len(argtypes), f"{name}: header has {len(params)} parameters, ctypes binds {len(argtypes)}"
        for i, (decl, ct) in enumerate(zip(params, argtypes)):
            if "*" in decl:
                want = ctypes.c_void_p
            elif decl.startswith(("int64_t", "long long")):
                want = ctypes.c_int64
            elif decl.startswith(("int32_t", "int ", "uint32_t")):
                want = ctypes.c_int32
            elif decl.startswith("float"):
                want = ctypes.c_float
            else:
                raise AssertionError(f"{name}: unrecognised parameter {decl!r}")
            assert ct is want or (want is ctypes.c_int32 and ct in (ctypes.c_int, ctypes.c_int32, ctypes.c_uint32)), \
                f"{name} argument {i} ({decl!r}): bound as {ct.__name__}"

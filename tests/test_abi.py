"""CPU-only: the C-ABI library builds, loads and exports every symbol include/sepb200.h declares."""
import ctypes
import os
import re

from multi_robot_slam_separators_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sepb200.h")).read()
    return sorted(set(re.findall(r"SEPB_API[^;(]*?\b(sepb_\w+)\s*\(", src)))


def test_library_builds_and_exports_all_symbols():
    path = build.build()
    L = ctypes.CDLL(path)
    names = declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(L, n), f"missing export {n}"


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return
    try:
        capi.Context(0)
    except capi.SepbError:
        return
    raise AssertionError("Context creation must fail loudly without a CUDA device")


def test_python_binding_structs_match_the_header(tmp_path):
    """The ctypes mirrors in capi.py against the C compiler's view of include/sepb200.h: size and every field offset
    (a silent drift here would shift arguments without any error)."""
    import subprocess
    structs = {"sepb_match_params": capi.MatchParams, "sepb_pnp_params": capi.PnpParams, "sepb_guess_params": capi.GuessParams,
               "sepb_verify_out": capi.VerifyOut, "sepb_ba_params": capi.BaParams}
    lines = []
    for cname, cls in structs.items():
        lines.append(f'printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, *_ in cls._fields_:
            lines.append(f'printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "sepb200.h"\nint main(void) {\n' + "\n".join(lines) +
                   "\nreturn 0;\n}\n")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])   # the header is plain C
    seen = 0
    for line in subprocess.check_output([str(exe)], text=True).splitlines():
        cname, field, value = line.split()
        cls = structs[cname]
        want = ctypes.sizeof(cls) if field == "size" else getattr(cls, field).offset
        assert int(value) == want, f"{cname}.{field}: header {value}, capi.py {want}"
        seen += 1
    assert seen == sum(len(c._fields_) + 1 for c in structs.values())
    # the status values the bindings and the host handlers rely on
    hdr = open(os.path.join(ROOT, "include", "sepb200.h")).read()
    for name, val in (("SEPB_PNP_OK", 1), ("SEPB_PNP_FEW_INLIERS", 3), ("SEPB_PNP_BA_FAILED", 4), ("SEPB_PNP_P3P", capi.PNP_P3P),
                      ("SEPB_PNP_AP3P", capi.PNP_AP3P), ("SEPB_NN_TENSOR", capi.NN_TENSOR)):
        assert re.search(rf"#define {name} \(?{val}\)?\b", hdr), name

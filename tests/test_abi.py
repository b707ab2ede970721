"""CPU checks of the drop-in boundary: libcharon_gpu.so loads, exports every symbol include/charon_gpu.h declares, the
ctypes structures agree with the C layout, and the product refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import numpy as np
import pytest
import charon_b200
from charon_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "charon_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cg_[a-z_0-9]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    lib = charon_b200.load_library()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(api.EXPORTS) == names           # the Python mirror lists exactly the header's functions


def test_struct_layouts_match_c_compiler(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "charon_gpu.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(cg_index_desc),sizeof(cg_build_desc),sizeof(cg_category_model),sizeof(cg_thresholds),sizeof(cg_batch),'
                   'sizeof(cg_counts_out),sizeof(cg_read_scalars),sizeof(cg_calls_out),sizeof(cg_info));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(t) for t in (api.IndexDesc, api.BuildDesc, api.CategoryModel, api.Thresholds, api.Batch, api.CountsOut,
                                  api.ReadScalars, api.CallsOut, api.Info)]
    assert got == want


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(charon_b200.CharonGpuError, match="no CUDA device"):
        charon_b200.GpuClassifier(k=19, w=41, bins=2, bin_size=1000, hash_funs=3, bin_to_cat=[0, 1], ncat=2, rows=np.zeros(1000, np.uint64))
    with pytest.raises(charon_b200.CharonGpuError):
        charon_b200.GpuClassifier.from_references([np.frombuffer(b"ACGT" * 100, np.uint8)], [0], [0], 1)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "charon_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in text and "oracle/" not in text and "liboracle" not in text, os.path.join(dirpath, f)

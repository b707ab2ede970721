"""The reference's built-in KDE datasets (src/dehost_main.cpp:23-206), read from csrc/kde_defaults.inc — the same table
the C++ host code compiles in — narrowed double->float like `std::vector<float>{...}` and sorted ascending as the
KDEParams constructor does (include/classify_stats.hpp:214-218)."""
import os
import re
import numpy as np

_INC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "kde_defaults.inc")


def _table(name):
    text = open(_INC).read()
    m = re.search(r"#define " + name + r" \\\n(.*?)\n\n", text + "\n\n", re.S)
    vals = [float(t) for t in re.findall(r"[-+0-9.eE]+", m.group(1).replace("\\", " "))]
    assert len(vals) == 500, (name, len(vals))
    return np.asarray(vals, dtype=np.float64).astype(np.float32)


def default_pos_sorted():
    return np.sort(_table("CHARON_KDE_DEFAULT_POS"))


def default_neg_sorted():
    return np.sort(_table("CHARON_KDE_DEFAULT_NEG"))

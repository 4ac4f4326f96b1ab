"""The documents cite evidence by file name: every profiles/ file, test, script and source file that DESIGN.md,
BASELINE.md, INTEGRATION.md, README.md, profiles/README.md and include/b200amg.h name must exist in the tree."""
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DOCS = ["DESIGN.md", "BASELINE.md", "INTEGRATION.md", "README.md", "profiles/README.md", "include/b200amg.h"]


def _expand(p):
    m = re.search(r"\{([^{}]*)\}", p)
    if not m:
        return [p]
    out = []
    for alt in m.group(1).split(","):
        out += _expand(p[:m.start()] + alt + p[m.end():])
    return out


@pytest.mark.parametrize("doc", DOCS)
def test_cited_files_exist(doc):
    text = open(os.path.join(ROOT, doc)).read()
    missing = []
    # profiles: `r02_name.json`, `profiles/r02_name*`, `r02_a_{1,2}gpu.json` (a cited stem may stand for its variants)
    for m in re.finditer(r"`((?:profiles/)?r0[12]_[A-Za-z0-9_{},.*\-]+)`", text):
        name = m.group(1) if m.group(1).startswith("profiles/") else "profiles/" + m.group(1)
        for p in _expand(name):
            if not glob.glob(os.path.join(ROOT, p)) and not glob.glob(os.path.join(ROOT, p) + "*"):
                missing.append(p)
    for m in re.finditer(r"(tests/[A-Za-z0-9_/]+\.py)(?:::([A-Za-z0-9_]+))?", text):
        path, test = m.group(1), m.group(2)
        if not os.path.exists(os.path.join(ROOT, path)):
            missing.append(path)
        elif test and ("def " + test) not in open(os.path.join(ROOT, path)).read():
            missing.append(path + "::" + test)
    for m in re.finditer(r"`((?:scripts|hypre_b200|oracle|include)/[A-Za-z0-9_/.]+\.(?:cu|cuh|hpp|c|h|py|sh|mk))`", text):
        if not os.path.exists(os.path.join(ROOT, m.group(1))):
            missing.append(m.group(1))
    assert not missing, missing

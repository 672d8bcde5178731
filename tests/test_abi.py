"""Every function declared in include/b200cnn.h is exported by one of the two native libraries (a prototype without
a definition would only show up when a C host links), and the libraries export nothing outside the b200_ namespace."""
import os
import re
import subprocess

from b200cnn.ops import _build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _exports(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
    return {l.split()[-1] for l in out.splitlines() if " T " in l}


def test_header_and_libraries_agree():
    hdr = open(os.path.join(ROOT, "include", "b200cnn.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) > 40
    cpu, cuda = _exports(_build.build_cpu()), _exports(_build.build_cuda())
    missing = sorted(d for d in declared if d not in cpu and d not in cuda)
    assert not missing, f"declared in b200cnn.h but defined nowhere: {missing}"
    stray = sorted(s for s in (cpu | cuda) if not s.startswith(("b200_", "_Z", "_init", "_fini", "__")))
    assert not stray, f"exported without the b200_ prefix: {stray}"
    assert not (cpu & cuda & declared), "a function is defined in both libraries"

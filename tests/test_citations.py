"""Every `file.ext:line` reference citation in the headers, the sources, the oracle and the two design documents must
name a file that exists in the reference tree and a line inside it.  Runs only where /root/reference is mounted."""
import collections
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def test_reference_citations_resolve():
    if not os.path.isdir(REF + "/src"):
        pytest.skip("reference tree not present here")
    index = collections.defaultdict(list)
    for root, _, files in os.walk(REF + "/src"):
        for f in files:
            index[f].append(os.path.join(root, f))
    for f in ("configure.ac", "README.md"):
        index[f].append(os.path.join(REF, f))
    lines = {}

    def n_lines(p):
        if p not in lines:
            with open(p, errors="ignore") as fh:
                lines[p] = sum(1 for _ in fh)
        return lines[p]

    own = {os.path.basename(p) for p in glob.glob(os.path.join(ROOT, "**", "*"), recursive=True)}
    pat = re.compile(r"([A-Za-z0-9_\-]+\.(?:cpp|hpp|h|c|ac|am)):(\d+)(?:-(\d+))?")
    sources = []
    for g in ("include/**/*.h*", "nodexa_chain_core_b200/**/*.py", "nodexa_chain_core_b200/csrc/*", "oracle/*.c", "oracle/*.h",
              "oracle/*.cpp", "oracle/*.py", "tests/*.py", "examples/**/*.c*", "DESIGN.md", "INTEGRATION.md"):
        sources += [p for p in glob.glob(os.path.join(ROOT, g), recursive=True) if os.path.isfile(p) and not p.endswith((".so", ".inc"))]
    checked, bad = 0, []
    for path in sources:
        for m in pat.finditer(open(path, errors="ignore").read()):
            name, last = m.group(1), int(m.group(3) or m.group(2))
            if name not in index:
                if name not in own:
                    bad.append((os.path.relpath(path, ROOT), m.group(0), "no such file in the reference"))
                continue
            checked += 1
            if not any(n_lines(p) >= last for p in index[name]):
                bad.append((os.path.relpath(path, ROOT), m.group(0), "line beyond the end of the file"))
    assert not bad, bad[:20]
    assert checked > 300

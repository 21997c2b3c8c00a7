#!/usr/bin/env python
"""Extracts the nanobind surface of the sparse path from the reference (quadrants/python/export_lang.cpp, the block from
"// Sparse Matrix" to "// Mesh Class", plus Program.create_sparse_matrix / make_sparse_matrix_from_ndarray) into
tests/golden/nanobind_surface.json.  Run in a container that has /root/reference; the JSON is committed so that the
surface test also runs where the reference is absent."""
import json
import os
import re

SRC = "/root/reference/quadrants/python/export_lang.cpp"
text = open(SRC).read()
block = text[text.index("// Sparse Matrix"):text.index("// Mesh Class")]
surface = {"classes": {}, "functions": sorted(set(re.findall(r'm\.def\("(\w+)"', block)))}
# macro-generated classes: expand the two macros by hand-written patterns
parts = re.split(r"(?=nb::class_<)|(?=#define )|(?=\n  m\.def\()|(?=\n  [A-Z_]+\()", block)
for part in parts:
    m = re.match(r'nb::class_<.*?>\(m, "(\w+)"\)', part, flags=re.S)
    if m:
        surface["classes"][m.group(1)] = sorted(set(re.findall(r'\.def\("(\w+)"', part)))
mac = re.search(r"#define MAKE_SPARSE_MATRIX.*?mat_vec_mul<[^\n]*\n", block, flags=re.S).group(0)
eigen_methods = sorted(set(re.findall(r'\.def\("(\w+)"', mac)))
for vt, st in (("f", "ColMajor"), ("f", "RowMajor"), ("d", "ColMajor"), ("d", "RowMajor")):
    surface["classes"][f"{vt}{st}_EigenSparseMatrix"] = eigen_methods
mac = re.search(r"#define REGISTER_EIGEN_SOLVER.*?info\);", block, flags=re.S).group(0)
solver_methods = sorted(set(re.findall(r'\.def\("(\w+)"', mac)))
for m in re.finditer(r"REGISTER_EIGEN_SOLVER\((\w+), (\w+), (\w+), \w\)", block):
    surface["classes"]["EigenSparseSolver" + "".join(m.groups())] = solver_methods
surface["program_methods"] = sorted(set(re.findall(r'\.def\(\s*"(create_sparse_matrix|make_sparse_matrix_from_ndarray)"', text)))
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "nanobind_surface.json")
json.dump(surface, open(out, "w"), indent=1, sort_keys=True)
print(out, {k: len(v) for k, v in surface["classes"].items()}, surface["functions"], surface["program_methods"])

"""Extract the golden vectors held by the reference's own unit specs into reference_vectors.json.

Run once in the build container (reads /root/reference, which does not exist on the
GPU box): python tests/golden/extract_golden.py
Sources: /root/reference/src/test/MeshManagerTests.cpp:180-228,
TriangleNodesProvisionerTests.cpp:133-161,289-378,420-463, QuadNodesProvisionerTests.cpp:65-193,
BlitzHelpersTests.cpp:99-210, JacobiBuildersTests.cpp.  Also copies the two small gmsh
inputs the parity tests run on (coarse_box.msh, coarse_box_quads.msh) as test fixtures.
"""
import json
import os
import re
import shutil

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def grab(path, name):
    """Return the numbers of the blitz comma-initialiser `name = a, b, c, ...;` (first occurrence)."""
    src = open(os.path.join(REF, path)).read()
    m = re.search(r"\b" + re.escape(name) + r"\s*=\s*([-+0-9.eE,\s]+?);", src)
    if not m:
        raise KeyError(name)
    return [float(t) for t in m.group(1).replace("\n", " ").split(",") if t.strip()]


def main():
    out = {}
    mm = "src/test/MeshManagerTests.cpp"
    for n in ["expectedVerts", "expectedBcTable", "expectedElements", "expectedEToE", "expectedEToF"]:
        out["mesh_" + n] = grab(mm, n)
    tt = "src/test/TriangleNodesProvisionerTests.cpp"
    for n in ["Lift_expected", "expectedDr", "expectedDs", "expectedx", "expectedy",
              "expectedvmapM", "expectedvmapP", "expectedvmapB", "expectedmapB", "expectedmapW"]:
        out["tri_" + n] = grab(tt, n)
    qt = "src/test/QuadNodesProvisionerTests.cpp"
    src = open(os.path.join(REF, qt)).read()
    for n in set(re.findall(r"\b(\w*[eE]xpected\w*)\s*=\s*[-+0-9.]", src)):
        try:
            out["quad_" + n] = grab(qt, n)
        except KeyError:
            pass
    with open(os.path.join(HERE, "reference_vectors.json"), "w") as fh:
        json.dump(out, fh, indent=0)
    for f in ["coarse_box.msh", "coarse_box_quads.msh"]:
        shutil.copy(os.path.join(REF, "input", f), os.path.join(HERE, f))
    print({k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()

"""The committed reference vectors ARE what the reference produces: where `/root/reference` exists (the build
container; not the GPU box, where this test skips), run tests/golden/make_reference_vectors.py -- the reference's own
`src.*` on the stand-ins of oracle/ref_shim -- into a scratch directory and compare every array and every JSON value
with the committed file.  A change of the stand-ins, of the generator or of a golden file that is not mirrored in
the others fails here (bit for bit in the container the files were made in; to float rounding on a host whose BLAS
rounds differently).  CPU only, a few seconds."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REFERENCE, "src")), reason="the reference tree is not on this machine")
def test_the_committed_vectors_regenerate_from_the_reference(tmp_path):
    env = dict(os.environ, OBST_GOLDEN_OUT=str(tmp_path))
    run = subprocess.run([sys.executable, os.path.join(GOLDEN, "make_reference_vectors.py")], env=env, cwd=ROOT,
                         capture_output=True, text=True, timeout=1500)
    assert run.returncode == 0, run.stdout[-2000:] + run.stderr[-2000:]
    assert "FAILED" not in run.stdout, run.stdout[-2000:]
    made = sorted(os.listdir(tmp_path))
    assert made == sorted(f for f in os.listdir(GOLDEN) if f.startswith("reference_")), made
    for name in made:
        if name.endswith(".json"):
            assert json.load(open(tmp_path / name)) == json.load(open(os.path.join(GOLDEN, name))), name
            continue
        with np.load(tmp_path / name) as new, np.load(os.path.join(GOLDEN, name)) as old:
            assert sorted(new.files) == sorted(old.files), name
            for key in old.files:
                a, b = old[key], new[key]
                assert a.dtype == b.dtype and a.shape == b.shape, (name, key)
                if np.array_equal(a, b, equal_nan=a.dtype.kind == "f"):  # what this container gives: bit for bit
                    continue
                # another host's BLAS may round differently: same numbers to float rounding, same decisions nearly all
                if a.dtype.kind == "f":
                    scale = float(np.linalg.norm(a.astype(np.float64))) + 1e-30
                    assert float(np.linalg.norm(a.astype(np.float64) - b.astype(np.float64))) <= 1e-4 * scale, (name, key)
                else:
                    assert np.mean(a == b) >= 0.98, (name, key)

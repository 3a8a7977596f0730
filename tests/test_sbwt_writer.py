"""The synthetic inputs of tests and bench.py ride on tools/fin_datagen's own plain-matrix SBWT writer; this pins it,
byte for byte, to the reference's in-memory constructor (oracle/_ref/sbwt_inmem = a 20-line driver around
SBWT/include/sbwt/NodeBOSSInMemoryConstructor.hh:98-188 + SBWT::serialize, SBWT.hh:395-424) for
k in {4, 12, 31, 63} and precalc tables of 0 / 2 / 8 characters."""
import os
import subprocess

import pytest

import helpers as H

SBWT_INMEM = os.path.join(H.REF_DIR, "sbwt_inmem")
DATAGEN = os.path.join(H.TOOLS, "fin_datagen")

pytestmark = pytest.mark.skipif(not os.path.exists(SBWT_INMEM), reason="oracle/_ref/sbwt_inmem is built where /root/reference exists")


def _datagen(*args, cwd):
    if not os.path.exists(DATAGEN):
        subprocess.run(["make", "-C", H.TOOLS], check=True, capture_output=True)
    r = subprocess.run([DATAGEN, *[str(a) for a in args]], cwd=cwd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def _unitigs(tmp, k):
    if k == 4:  # the paper example: tiny graph with dummy-node chains of every length
        p = os.path.join(H.GOLDEN, "example_fna_k4", "unitigs.fna")
        return p
    glen = {12: 4000, 31: 30000, 63: 20000}[k]
    _datagen("genome", "g.fna", glen, 100 + k, 3, 3, max(2 * k, glen // 25), 0.02, cwd=tmp)
    _datagen("unitigs", "g.fna", k, "u.fna", 200 + k, cwd=tmp)
    return os.path.join(tmp, "u.fna")


@pytest.mark.parametrize("k", [4, 12, 31, 63])
@pytest.mark.parametrize("precalc", [0, 2, 8])
def test_datagen_sbwt_equals_reference_constructor(k, precalc, tmp_path):
    tmp = str(tmp_path)
    u = _unitigs(tmp, k)
    _datagen("sbwt", u, k, "ours.sbwt", precalc, cwd=tmp)
    r = subprocess.run([SBWT_INMEM, u, os.path.join(tmp, "ref.sbwt"), str(k), str(precalc)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert H.files_equal(os.path.join(tmp, "ours.sbwt"), os.path.join(tmp, "ref.sbwt"))

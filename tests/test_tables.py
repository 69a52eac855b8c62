"""The marching-cubes tables: generated from Bloomenthal's construction (tools/gen_mc_tables.py), pinned by SHA-256, and - where
the reference is present - diffed against its g_triTableCache (Parsip100/PS_SimdPoly/include/_CellConfigTable.h:60-317)."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHA = "0be1686c5a220a0b144e0bfeae245fbf2d4814d5d9dbd868ddb223c78e29a13c"


def test_committed_header_carries_the_pinned_hash_and_regenerates_identically(tmp_path):
    hdr = open(os.path.join(ROOT, "parsip_b200", "csrc", "mc_tables.h")).read()
    assert "sha256 = " + SHA in hdr
    out = tmp_path / "mc_tables.h"
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_mc_tables.py"), "--out", str(out)])
    assert open(out).read() == hdr, "parsip_b200/csrc/mc_tables.h is not what tools/gen_mc_tables.py generates"


@pytest.mark.skipif(not os.path.isdir("/root/reference/Parsip100"), reason="reference tree not present on this box")
def test_generated_table_equals_the_reference_table():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_mc_tables.py"), "--check", "/root/reference", "--out", os.devnull],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    shas = re.findall(r"sha256 (?:generated|reference): ([0-9a-f]{64})", r.stdout)
    assert shas == [SHA, SHA], r.stdout

"""-m gpu: __graft_entry__.smoke() is what the driver runs on the GPU box before the bench; keep it under test."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_smoke_entry_point(ctx, capsys):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import __graft_entry__ as g
    g.smoke()
    assert "smoke ok" in capsys.readouterr().out

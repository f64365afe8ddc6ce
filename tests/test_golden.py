"""The travelling restatement oracle/ref_heads.py reproduces the committed golden vectors that were
generated from the reference's own code (tests/golden/make_golden.py).  CPU only."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden  # noqa: E402

from oracle import ref_heads  # noqa: E402


@pytest.mark.parametrize("name", sorted(make_golden.CASES))
def test_ref_heads_reproduces_reference_fixtures(name):
    inputs, kw = make_golden.case(name)
    out = make_golden.flatten(ref_heads.head_step(**inputs, **kw))
    gold = np.load(os.path.join(os.path.dirname(make_golden.__file__), "heads_%s.npz" % name))
    assert set(gold.files) == set(out)
    for key in gold.files:
        g, o = gold[key], out[key]
        if g.dtype.kind in "iu" or key.endswith(("gt_boxes", "pgt_boxes", "gt_weights")):
            assert np.array_equal(g, o), key
        else:
            np.testing.assert_allclose(o, g, rtol=1e-6, atol=1e-7, err_msg=key)

"""Dev-container-only cross-checks of the oracle's constant tables against the
reference SOURCE TEXT (skipped wherever /root/reference is absent, e.g. on the
GPU box).  They guard against transcription slips in the rotation constants
(src/transform/forward.rs) and the sub-pel filter bank (src/mc.rs:70-179)."""
import collections
import os
import re

import pytest

REF = "/root/reference/src"
HERE = os.path.dirname(os.path.abspath(__file__))
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")


def _strip_comments(text):
    return "\n".join(l.split("//")[0] for l in text.splitlines())


def test_rotation_constants_multiset():
    rust = _strip_comments(open(os.path.join(REF, "transform/forward.rs")).read())
    rust = rust[rust.index("fn daala_fdct_ii_2_asym"):rust.index("fn fidentity4")]
    # every (mult, shift) literal pair handed to tx_mul via a kernel call
    ref_pairs = collections.Counter(
        (int(a), int(b)) for a, b in re.findall(r"\((\d+),\s*(\d+)\)", rust))
    for i, j in enumerate([0, 8, 4, 12, 2, 10, 6, 14, 1, 9, 5, 13, 3, 11, 7, 15]):
        ref_pairs[(i, j)] -= 1  # the reorder_4(i, j) calls of daala_fdct64 are not constants
    ref_pairs = +ref_pairs
    c = open(os.path.join(HERE, "..", "oracle", "r1o_tx.c")).read()
    c = re.sub(r"/\*.*?\*/", "", c, flags=re.S)
    c = c[c.index("fdct_ii_2_asym"):c.index("1-D dispatch") if "1-D dispatch" in c else len(c)]
    mine = collections.Counter()
    for m in re.finditer(r"\b(rot_half|rot_neg|rot_pi4|rot|mul)\(([^;]*?)\);", c):
        nums = [int(x) for x in re.findall(r"(?<![\w\[])(\d+)(?![\w\]])", m.group(2))]
        if m.group(1) == "rot_neg":
            nums = nums[1:]  # leading avg flag
        if m.group(1) == "mul":
            nums = nums[-2:]
        assert len(nums) % 2 == 0, m.group(0)
        for k in range(0, len(nums), 2):
            mine[(nums[k], nums[k + 1])] += 1
    tab = re.search(r"ST0_32\[16\]\[6\] = \{(.*?)\};", c, flags=re.S).group(1)
    nums = [int(x) for x in re.findall(r"\d+", tab)]
    for k in range(0, len(nums), 2):
        mine[(nums[k], nums[k + 1])] += 1
    assert mine == ref_pairs, (mine - ref_pairs, ref_pairs - mine)


def test_subpel_filter_bank():
    rust = _strip_comments(open(os.path.join(REF, "mc.rs")).read())
    body = rust[rust.index("const SUBPEL_FILTERS"):rust.index("pub(crate) mod native")]
    body = body[body.index("= ["):]
    ref = [int(x) for x in re.findall(r"-?\d+", body)]
    assert len(ref) == 6 * 16 * 8
    c = open(os.path.join(HERE, "..", "oracle", "r1o_mc.c")).read()
    tab = re.search(r"FILT\[6\]\[16\]\[8\] = \{(.*?)\n\};", c, flags=re.S).group(1)
    tab = re.sub(r"/\*.*?\*/", "", tab, flags=re.S)
    mine = [int(x) for x in re.findall(r"-?\d+", tab)]
    assert mine == ref

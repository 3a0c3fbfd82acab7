"""The evidence under profiles/ is produced by tools/*.py; these checks run them on the CPU and pin the claims the
documentation makes from their output: which hardware instructions each kernel family contains (SASS of the built
library) and what a ResNet-50 / BERT-large step consists of (committed launch lists)."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tony_b200", "lib", "libtony_b200_kernels.so")


def _run(*argv):
    r = subprocess.run([sys.executable, *argv], capture_output=True, text=True, cwd=ROOT, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout


@pytest.mark.skipif(shutil.which("cuobjdump") is None or not os.path.exists(SO), reason="needs cuobjdump and the built library")
def test_sass_table_shows_the_blackwell_paths(tmp_path):
    out = str(tmp_path / "sass.md")
    _run("tools/sass_table.py", "--so", SO, "--out", out)
    rows = {}
    hdr = None
    for line in open(out):
        if line.startswith("| kernel |"):
            hdr = [c.strip() for c in line.strip().strip("|").split("|")]
        elif line.startswith("| `") and hdr:
            cells = [c.strip() for c in line.strip().strip("|").split("|")]
            rows[cells[0].strip("`")] = {h: (int(c) if c.isdigit() else 0) for h, c in zip(hdr[1:], cells[1:])}
    assert rows, "no kernel rows parsed"

    def fam(prefix):
        got = {k: v for k, v in rows.items() if prefix in k}
        assert got, f"no kernel matching {prefix}"
        return got

    for k, v in {**fam("gemm_bf16_tcgen05"), **fam("conv_bf16_tcgen05_kernel"), **fam("conv3x3_halo")}.items():
        assert v["UTCHMMA"] > 0 and v["UTMALDG"] > 0 and v["LDTM"] > 0 and v["UTCBAR"] > 0, (k, v)   # tcgen05 + TMA + TMEM
    for k, v in fam("gemm_bf16_tcgen05").items():
        assert v["UTMASTG"] > 0, (k, v)                                                   # TMA-store epilogue
    for k, v in fam("attention_").items():
        if "delta" not in k:
            assert v["UTCHMMA"] > 0 and v["UTMALDG"] > 0, (k, v)
    assert all(v["LDGMC"] > 0 for v in fam("fused_reduce_optim_nvls").values())           # multimem.ld_reduce
    assert all(v.get("HMMA", 0) == 0 for v in rows.values()), "legacy mma.sync in the library"


@pytest.mark.parametrize("csv_name,own_share,forbidden", [
    ("r2_c66_launches_resnet50_n1_warm.csv", 95.0, ("cudnn", "cutlass", "xmma")),
    ("r2_c66_launches_bert_large_n1_warm.csv", 80.0, ("cutlass", "xmma", "ampere")),
])
def test_committed_launch_lists_match_the_claims(csv_name, own_share, forbidden):
    path = os.path.join(ROOT, "profiles", csv_name)
    if not os.path.exists(path):
        pytest.skip("launch list not in this checkout")
    text = _run("tools/launch_summary.py", path)
    first = text.splitlines()[0]
    share = float(first.split("own-kernel share")[1].split("%")[0])
    assert share >= own_share, first
    low = text.lower()
    for word in forbidden:
        assert word not in low, f"{word} kernel in {csv_name}"

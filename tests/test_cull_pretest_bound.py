"""The cull kernel's rejection pass (csrc/proj.cu: make_precam / surely_culled) may only reject what the reference's
full test rejects.  A float32 numpy restatement of it (tests/cull_pretest_check.py) is checked against the oracle's
keep mask on the three bench scene shapes plus adversarial Gaussians (next to a camera, on the frustum edges, huge /
anisotropic).  The kernel itself is pinned on the GPU: its masks must equal the oracle's bit for bit
(tests/test_gpu_forward.py, tests/test_gpu_vs_reference_cuda.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def test_rejection_pass_never_rejects_a_kept_pair(monkeypatch, capsys):
    import cull_pretest_check as C
    monkeypatch.setattr(sys, "argv", ["cull_pretest_check.py", "15000"])
    assert C.main() == 0
    out = capsys.readouterr().out
    assert "TOTAL WRONG 0" in out
    # and it must stay effective where it matters: >= 99 % of the rejectable pairs of the aerial scene
    line = next(l for l in out.splitlines() if l.startswith("c4_bigcity_102m  "))
    assert float(line.split("(")[1].split("%")[0]) >= 99.0, line

"""CPU: host-side pieces of bench.py — the clock sampler's bookkeeping, the algorithmic-bytes table of the newest
kernels, and the reference arm's JSON line (the oracle port timed on the host cores)."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


class _Proc:
    def terminate(self):
        pass


class _Thread:
    def join(self, timeout=None):
        pass


def test_clock_sampler_reports_only_samples_after_mark():
    s = bench.ClockSampler(0)
    s.proc, s.t = _Proc(), _Thread()
    s.lines = ["1200, 1965, 300.0, Not Active, Not Active, Not Active, Not Active\n"] * 3  # start-up / warm-up samples
    s.mark()
    s.lines += ["1965, 1965, 900.0, Not Active, Not Active, Not Active, Active\n",
                "1950, 1965, 910.0, Not Active, Not Active, Not Active, Not Active\n",
                "garbage\n"]
    out = s.stop()
    assert out["samples"] == 2 and out["sm_mhz"] == 1957.5 and out["sm_max_mhz"] == 1965.0
    assert out["reasons"] == ["sw_power_cap"]


def test_clock_sampler_without_nvidia_smi():
    s = bench.ClockSampler(0)
    assert s.stop()["reasons"] == ["nvidia-smi unavailable"]


def test_alg_bytes_of_the_classifier_glue():
    M, Na, Nb = 1000, 512, 131
    fwd = [0, M, Na, Na, 0, Nb, Nb, 0.7, 1, 0, 0]
    bwd = [0, 0, M, Na, Nb, 0.7, 1, 0, Na, 0, Nb, 0]
    assert bench._alg_bytes("pcs_relu_concat_dropout_forward", fwd) == 8 * M * (Na + Nb)
    assert bench._alg_bytes("pcs_relu_concat_dropout_backward", bwd) == 4 * M * (3 * Na + 2 * Nb)
    assert bench._alg_bytes("pcs_not_a_kernel", []) == 0


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["unit"] == "points/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_fill_to_budget_balances_ranks():
    """The loader's balancing step (dist.fill_to_budget): n of the candidates, total as close to the budget as single
    swaps get; batches of different ranks then differ by a handful of points instead of ~2 %."""
    import numpy as np

    from pointcloudsegmentation_b200.dist import fill_to_budget
    from pointcloudsegmentation_b200.synthetic import s3dis_batch_to_budget

    rng = np.random.default_rng(0)
    for seed in range(5):
        sizes = rng.integers(6144, 10001, 24).tolist()
        keep = fill_to_budget(sizes, 16, 16 * 8072)
        assert len(keep) == 16 == len(set(keep)) and keep == sorted(keep)
        assert abs(sum(sizes[i] for i in keep) - 16 * 8072) <= 16
    assert fill_to_budget([5, 6, 7], 3, 100) == [0, 1, 2]  # nothing to swap with
    d = s3dis_batch_to_budget(7, 4, 4 * 8072)
    assert abs(d["xyzs"].shape[0] - 4 * 8072) <= 64
    assert d["block_ids"].min() == 0 and d["block_ids"].max() == 3 and np.all(np.diff(d["block_ids"]) >= 0)
    assert d["feats"].shape[0] == d["labels"].shape[0] == d["xyzs"].shape[0]

"""Short chaos run on the CPU plane: peers are spawned and SIGKILLed at random while the survivors keep reducing
(reference: python/tests/stress_tests/basic_stress_test/stresstest_orchestrator.py, 8 h manual run; here 20 s)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent / "stress"))
from orchestrator import run  # noqa: E402


def test_random_spawn_and_kill_never_stalls():
    out = run(seconds=20, max_peers=4, stall=12, device="cpu", seed=3, min_kills=1)
    assert not out["stalled"], out
    assert not out.get("fatal"), out
    assert out["reductions"] > 50, out
    assert out["killed"] >= 1, out


def test_cross_step_async_reductions_under_churn():
    """The reference's second stress test (python/tests/stress_tests/cross_step_async_reduces_test): ten tensors per step
    through all_reduce_multiple_with_retry on a thread that overlaps the next step, peers coming and going."""
    out = run(seconds=20, max_peers=4, stall=12, device="cpu", seed=5, mode="cross_step", min_kills=1)
    assert not out["stalled"], out
    assert not out.get("fatal"), out
    assert out["reductions"] > 50, out
    assert out["killed"] >= 1, out

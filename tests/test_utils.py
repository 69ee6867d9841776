import time

from arbius_b200.utils import LatencyHistogram, StageTimer


def test_stage_timer_cpu_fallback_and_sums_repeated_stages():
    t = StageTimer(device="cpu")
    for _ in range(2):
        with t.stage("a"):
            time.sleep(0.01)
    with t.stage("b"):
        pass
    s = t.summary()
    assert s["a"] >= 15.0 and 0.0 <= s["b"] < 5.0
    t.reset()
    assert t.summary() == {}


def test_latency_histogram_quantiles_and_prometheus():
    h = LatencyHistogram()
    for ms in [100.0] * 50 + [130.0] * 40 + [2000.0] * 10:
        h.observe(ms)
    assert h.n == 100 and abs(h.total - (5000 + 5200 + 20000)) < 1e-6
    assert 79.0 <= h.quantile(0.5) <= 101.0 and 1500.0 <= h.quantile(0.99) <= 2520.0
    text = h.prometheus("arbius_task_latency_ms", 'model="anythingv3"')
    assert 'le="+Inf"} 100' in text and "arbius_task_latency_ms_count{model=\"anythingv3\"} 100" in text
    assert LatencyHistogram().quantile(0.5) == 0.0

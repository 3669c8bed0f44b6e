"""The headline numbers quoted in README.md / BASELINE.md / DESIGN.md are the ones in the bench lines committed under
profiles/ (round 1's README advertised a scaling figure no committed file held)."""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        return json.loads(f.read().strip().splitlines()[-1])


def _text(name):
    with open(os.path.join(ROOT, name)) as f:
        return f.read()


def test_headline_numbers_match_the_committed_bench_lines():
    n1, n2, n4, n8 = (_line("r02n_bench_n1.json"), _line("r02o_bench_n2.json"), _line("r02n_bench_n4.json"),
                      _line("r02o_bench_n8.json"))
    m = lambda d: "%.1f M" % (d["value"] / 1e6)
    for doc in ("README.md", "BASELINE.md", "DESIGN.md"):
        t = _text(doc)
        for d in (n1, n2, n4, n8):
            assert m(d) in t, (doc, m(d))
        assert "%.2f" % (n8["value"] / n1["value"]) in t, doc          # the N=8 ratio
    assert "%.1f M" % (n1["e2e"]["value"] / 1e6) in _text("BASELINE.md")
    assert "%.2f" % n1["roofline"]["frac"] in _text("BASELINE.md")
    for d, n in ((n1, 1), (n2, 2), (n4, 4), (n8, 8)):
        assert d["n_gpus"] == n and d["metric"] == "deepfm_train_samples_per_sec" and d["clocks"]["reasons"] == []

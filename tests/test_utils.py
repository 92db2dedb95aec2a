"""Host-side helpers: the section profiler (reference python/examples/nanogpt_diloco/profiler.py) and NUMA binding."""
import json
import time

import pytest

from pccl_b200.utils import Profiler, ProfilerCollection, bind_to_device


def busy(ms):
    end = time.perf_counter() + ms / 1e3
    while time.perf_counter() < end:
        pass


def profile_one_step(p):
    with p.section("outer step"):
        with p.section("inner"):
            for _ in range(3):
                with p.section("forward"):
                    busy(2)
                with p.section("backward"):
                    busy(3)
        with p.session("reduce"):  # the reference's spelling
            busy(4)


def test_report_aggregates_and_nests():
    p = Profiler()
    profile_one_step(p)
    inner = p.root.children["outer step"].children["inner"]
    assert inner.children["forward"].count == 3 and inner.children["backward"].count == 3
    assert 0.006 <= inner.children["forward"].total < 0.05
    assert p.root.children["outer step"].total >= inner.total + p.root.children["outer step"].children["reduce"].total
    lines = p.report().splitlines()
    assert lines[0].startswith("outer step:") and lines[1].startswith("  inner:") and lines[2].startswith("    forward:") and "3x" in lines[2]
    assert len(p.intervals) == 9
    # an exception inside a section closes it
    with pytest.raises(ValueError):
        with p.section("broken"):
            raise ValueError
    assert p.root.children["broken"].count == 1 and len(p._stack) == 1


def test_trace_file_and_pictures(tmp_path):
    p = Profiler()
    profile_one_step(p)
    trace = tmp_path / "trace.json"
    p.export_chrome_trace(str(trace))
    events = [e for e in json.loads(trace.read_text())["traceEvents"] if e["ph"] == "X"]
    assert len(events) == 9 and events[0]["name"] == "outer step" and events[0]["ts"] == 0
    by_name = {}
    for e in events:
        by_name.setdefault(e["name"], []).append(e)
    outer = by_name["outer step"][0]
    assert all(outer["ts"] <= e["ts"] and e["ts"] + e["dur"] <= outer["ts"] + outer["dur"] + 1 for e in events)  # children lie inside
    svg = tmp_path / "timeline.svg"
    p.export_timeline(str(svg), width=1200, row_height=30)
    text = svg.read_text()
    assert text.count("<rect") == 10 and "forward" in text
    pil = pytest.importorskip("PIL")
    png = tmp_path / "timeline.png"
    p.export_timeline(str(png), width=1200, row_height=30)
    from PIL import Image
    img = Image.open(png)
    assert img.size == (1200, 90)  # three nesting levels
    assert len(img.getcolors(maxcolors=1 << 20)) > 4  # bars of several colours, not a blank sheet
    # one frame per step, same time scale
    coll = ProfilerCollection()
    for step in range(3):
        q = Profiler()
        profile_one_step(q)
        coll.add_profiler(q, f"step {step}")
    gif = tmp_path / "steps.gif"
    coll.render_as_gif(str(gif), fps=4, width=800, row_height=24)
    anim = Image.open(gif)
    assert anim.n_frames == 3
    assert Profiler().export_timeline(str(tmp_path / "empty.png")) is None


def test_numa_binding_reports_what_it_did():
    got = bind_to_device(0)
    assert isinstance(got["bound"], bool) and (got["bound"] or got["reason"])  # without a GPU: says why, changes nothing

"""Nested wall-clock sections: a tree report, a timeline picture and a trace file.

Counterpart of the reference's python/examples/nanogpt_diloco/profiler.py (sessions :8-56, timeline image :58-183,
`ProfilerCollection` video :216-278).  Differences: sections of the same name under the same parent are aggregated in the
report (count, total, average) while every single interval is still kept for the timeline; sections can bracket GPU work
(`sync_cuda=True` synchronises the device on both sides); besides the picture (PNG through Pillow, SVG when Pillow is not
installed) the intervals are written in the Chrome trace-event format, which chrome://tracing and Perfetto open; a collection
of profilers (one per outer step, say) becomes an animated GIF instead of an MP4, for which this image has no encoder."""
from __future__ import annotations

import colorsys
import json
import time
from contextlib import contextmanager
from typing import Dict, List, Optional, Tuple


class Section:
    def __init__(self, name: str):
        self.name = name
        self.total = 0.0
        self.count = 0
        self.children: Dict[str, "Section"] = {}

    def to_dict(self) -> dict:
        return {"name": self.name, "total_s": self.total, "count": self.count, "children": [c.to_dict() for c in self.children.values()]}


def _colour(name: str) -> Tuple[int, int, int]:
    """A pastel colour that depends on the name only, so a section keeps its colour from picture to picture."""
    h = 0
    for ch in name:
        h = (h * 131 + ord(ch)) % 3600
    r, g, b = colorsys.hls_to_rgb(h / 3600.0, 0.78, 0.55)
    return int(r * 255), int(g * 255), int(b * 255)


class Profiler:
    MAX_INTERVALS = 200_000  # the timeline stops growing there; the report keeps counting

    def __init__(self, sync_cuda: bool = False):
        self.root = Section("root")
        self._stack: List[Section] = [self.root]
        self.sync_cuda = sync_cuda
        self.intervals: List[Tuple[str, int, float, float]] = []  # name, depth, start, end (perf_counter seconds)

    @contextmanager
    def section(self, name: str):
        parent = self._stack[-1]
        node = parent.children.setdefault(name, Section(name))
        depth = len(self._stack) - 1
        self._stack.append(node)
        self._sync()
        t0 = time.perf_counter()
        try:
            yield node
        finally:
            self._sync()
            t1 = time.perf_counter()
            node.total += t1 - t0
            node.count += 1
            if len(self.intervals) < self.MAX_INTERVALS:
                self.intervals.append((name, depth, t0, t1))
            self._stack.pop()

    session = section  # the reference's name for it

    def _sync(self) -> None:
        if self.sync_cuda:
            import torch
            if torch.cuda.is_available():
                torch.cuda.synchronize()

    # ---- text ---------------------------------------------------------------------------------------------------------
    def report(self, node: Optional[Section] = None, indent: int = 0) -> str:
        node = node or self.root
        lines = []
        if node is not self.root:
            avg = node.total / max(1, node.count)
            lines.append(f"{'  ' * indent}{node.name}: {node.total * 1e3:.2f} ms total, {node.count}x, {avg * 1e3:.3f} ms avg")
        for c in node.children.values():
            lines.append(self.report(c, indent + (0 if node is self.root else 1)))
        return "\n".join(l for l in lines if l)

    def print_report(self) -> None:
        print(self.report())

    def dump(self, path: str) -> None:
        with open(path, "w") as f:
            json.dump(self.root.to_dict(), f, indent=1)

    # ---- trace file ---------------------------------------------------------------------------------------------------
    def export_chrome_trace(self, path: str, process_name: str = "pccl_b200", pid: int = 0) -> None:
        """Complete events ("ph":"X", microseconds) on one track; nesting follows from the timestamps."""
        if not self.intervals:
            events = []
        else:
            origin = min(t0 for _, _, t0, _ in self.intervals)
            events = [{"name": name, "ph": "X", "pid": pid, "tid": 0, "ts": (t0 - origin) * 1e6, "dur": (t1 - t0) * 1e6, "args": {"depth": depth}}
                      for name, depth, t0, t1 in sorted(self.intervals, key=lambda iv: (iv[2], iv[1]))]
        events.insert(0, {"name": "process_name", "ph": "M", "pid": pid, "args": {"name": process_name}})
        with open(path, "w") as f:
            json.dump({"traceEvents": events, "displayTimeUnit": "ms"}, f)

    # ---- picture ------------------------------------------------------------------------------------------------------
    def _layout(self, width: int, row_height: int, span: Optional[Tuple[float, float]] = None):
        """Rectangles (x0, y0, x1, y1, name, seconds) of all intervals, one row per nesting depth."""
        if not self.intervals:
            return [], 0
        lo = span[0] if span else min(t0 for _, _, t0, _ in self.intervals)
        hi = span[1] if span else max(t1 for _, _, _, t1 in self.intervals)
        scale = (width - 1) / max(hi - lo, 1e-9)
        rects = []
        rows = 0
        for name, depth, t0, t1 in self.intervals:
            x0, x1 = (t0 - lo) * scale, (t1 - lo) * scale
            rects.append((x0, depth * row_height, max(x1, x0 + 1), (depth + 1) * row_height - 2, name, t1 - t0))
            rows = max(rows, depth + 1)
        return rects, rows * row_height

    def export_timeline(self, filename: Optional[str] = None, width: int = 2400, row_height: int = 40, return_image: bool = False,
                        span: Optional[Tuple[float, float]] = None):
        """Draws every recorded interval as a bar (x = time, row = nesting depth, label = name and milliseconds where it
        fits).  `filename` ending in .svg, or a missing Pillow, gives an SVG; otherwise whatever format Pillow derives from
        the name.  `return_image=True` returns the Pillow image instead of writing a file."""
        rects, height = self._layout(width, row_height, span)
        if not rects:
            return None
        want_svg = filename is not None and filename.lower().endswith(".svg")
        image_mod = None
        if not want_svg:
            try:
                from PIL import Image, ImageDraw, ImageFont
                image_mod = Image
            except ImportError:
                if return_image:
                    raise
                want_svg = True
                filename = (filename or "timeline") + ".svg"
        if want_svg:
            parts = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}" font-family="monospace" font-size="{row_height // 3}">',
                     f'<rect width="{width}" height="{height}" fill="white"/>']
            for x0, y0, x1, y1, name, secs in rects:
                r, g, b = _colour(name)
                parts.append(f'<rect x="{x0:.1f}" y="{y0}" width="{x1 - x0:.1f}" height="{y1 - y0}" fill="rgb({r},{g},{b})" stroke="black" stroke-width="0.5">'
                             f'<title>{name}: {secs * 1e3:.3f} ms</title></rect>')
                label = f"{name} {secs * 1e3:.2f} ms"
                if (x1 - x0) > len(label) * row_height * 0.21:
                    parts.append(f'<text x="{x0 + 3:.1f}" y="{y0 + row_height * 0.6:.1f}">{label}</text>')
            parts.append("</svg>")
            with open(filename, "w") as f:
                f.write("\n".join(parts))
            return None
        img = image_mod.new("RGB", (width, height), "white")
        draw = ImageDraw.Draw(img)
        font = ImageFont.load_default()
        for x0, y0, x1, y1, name, secs in rects:
            draw.rectangle([x0, y0, x1, y1], fill=_colour(name), outline=(0, 0, 0))
            label = f"{name} {secs * 1e3:.2f} ms"
            while label and draw.textlength(label, font=font) > (x1 - x0) - 6:  # shorten until it fits, drop when nothing is left
                label = label[:-1]
            if len(label) >= 3:
                draw.text((x0 + 3, y0 + (y1 - y0) / 2 - 5), label, fill=(0, 0, 0), font=font)
        if return_image:
            return img
        if filename:
            img.save(filename)
        return None


class ProfilerCollection:
    """Several profilers (one per outer step, one per peer, ...) drawn to the same time scale, as frames of one animation."""

    def __init__(self):
        self.profilers: List[Tuple[str, Profiler]] = []

    def add_profiler(self, profiler: Profiler, label: Optional[str] = None) -> None:
        self.profilers.append((label or f"profiler {len(self.profilers)}", profiler))

    def render_frames(self, width: int = 1600, row_height: int = 32):
        from PIL import Image, ImageDraw
        longest = max((max(t1 for _, _, _, t1 in p.intervals) - min(t0 for _, _, t0, _ in p.intervals)) for _, p in self.profilers if p.intervals)
        frames = []
        for label, p in self.profilers:
            if not p.intervals:
                continue
            start = min(t0 for _, _, t0, _ in p.intervals)
            bars = p.export_timeline(width=width, row_height=row_height, return_image=True, span=(start, start + longest))
            frame = Image.new("RGB", (width, bars.height + 20), "white")
            ImageDraw.Draw(frame).text((4, 4), f"{label}   (full width = {longest * 1e3:.1f} ms)", fill=(0, 0, 0))
            frame.paste(bars, (0, 20))
            frames.append(frame)
        height = max(f.height for f in frames)
        return [f if f.height == height else _pad(f, height) for f in frames]

    def render_as_gif(self, out_filename: str = "profiler.gif", fps: float = 2.0, width: int = 1600, row_height: int = 32) -> None:
        frames = self.render_frames(width, row_height)
        frames[0].save(out_filename, save_all=True, append_images=frames[1:], duration=int(1000 / fps), loop=0)


def _pad(frame, height):
    from PIL import Image
    out = Image.new("RGB", (frame.width, height), "white")
    out.paste(frame, (0, 0))
    return out

"""Nested wall-clock sections with a tree report (reference: python/examples/nanogpt_diloco/profiler.py:8-120;
the PIL timeline / MP4 export of the reference is not reproduced, the tree report and a JSON dump are)."""
from __future__ import annotations

import json
import time
from contextlib import contextmanager
from typing import Dict, List, Optional


class Section:
    def __init__(self, name: str):
        self.name = name
        self.total = 0.0
        self.count = 0
        self.children: Dict[str, "Section"] = {}

    def to_dict(self) -> dict:
        return {"name": self.name, "total_s": self.total, "count": self.count, "children": [c.to_dict() for c in self.children.values()]}


class Profiler:
    def __init__(self, sync_cuda: bool = False):
        self.root = Section("root")
        self._stack: List[Section] = [self.root]
        self.sync_cuda = sync_cuda

    @contextmanager
    def section(self, name: str):
        parent = self._stack[-1]
        node = parent.children.setdefault(name, Section(name))
        self._stack.append(node)
        self._sync()
        t0 = time.perf_counter()
        try:
            yield node
        finally:
            self._sync()
            node.total += time.perf_counter() - t0
            node.count += 1
            self._stack.pop()

    def _sync(self) -> None:
        if self.sync_cuda:
            import torch
            if torch.cuda.is_available():
                torch.cuda.synchronize()

    def report(self, node: Optional[Section] = None, indent: int = 0) -> str:
        node = node or self.root
        lines = []
        if node is not self.root:
            avg = node.total / max(1, node.count)
            lines.append(f"{'  ' * indent}{node.name}: {node.total * 1e3:.2f} ms total, {node.count}x, {avg * 1e3:.3f} ms avg")
        for c in node.children.values():
            lines.append(self.report(c, indent + (0 if node is self.root else 1)))
        return "\n".join(l for l in lines if l)

    def dump(self, path: str) -> None:
        with open(path, "w") as f:
            json.dump(self.root.to_dict(), f, indent=1)

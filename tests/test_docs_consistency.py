"""Keeps the written inventory honest: every environment variable the code reads is documented, and every file and test
that COVERAGE.md names exists."""
import re
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent


def test_every_environment_variable_is_documented():
    used = set()
    for path in list((REPO / "csrc").rglob("*.[ch]pp")) + list((REPO / "csrc").rglob("*.cu")):
        used |= set(re.findall(r'(?:env_u64|env_flag|getenv)\("(PCCL_[A-Z0-9_]+)"', path.read_text()))
    documented = set(re.findall(r"`(PCCL_[A-Z0-9_]+)`", (REPO / "INTEGRATION.md").read_text()))
    assert used - documented == set(), f"undocumented: {sorted(used - documented)}"


def test_coverage_md_points_at_things_that_exist():
    text = (REPO / "COVERAGE.md").read_text()
    cpp_cases = {}
    for src in (REPO / "tests" / "cpp").glob("test_*.cpp"):
        cpp_cases[src.stem] = set(re.findall(r"^TEST\((\w+)\)", src.read_text(), flags=re.M))
    all_cases = set().union(*cpp_cases.values())
    missing = []
    # files: anything in backticks that looks like a repository path
    for token in set(re.findall(r"`([A-Za-z0-9_./{},-]+)`", text)):
        if "/" not in token or token.startswith(("cpp:", "py:", "gpu:")) or token.startswith("/"):
            continue
        if not token.startswith(("csrc/", "include/", "pccl_b200/", "tests/", "tools/", "examples/", "profiles/", "docs/", "baseline/")):
            continue
        m = re.match(r"(.*)\.\{([^}]*)\}$", token)
        candidates = [f"{m.group(1)}.{ext}" for ext in m.group(2).split(",")] if m else [token]
        for c in candidates:
            if not (REPO / c).exists():
                missing.append(c)
    # C++ test cases: `cpp:binary/case`
    for binary, case in re.findall(r"`cpp:(test_\w+)/(\w+)`", text):
        if binary not in cpp_cases or (case not in cpp_cases[binary] and not any(c.startswith(case.rstrip("_")) for c in cpp_cases[binary])):
            missing.append(f"cpp:{binary}/{case}")
    # bare case names in backticks that follow a cpp: reference in the same cell
    for case in re.findall(r"`([a-z][a-z0-9_]{12,})`", text):
        if case in all_cases or "/" in case or "." in case:
            continue
        if re.match(r"(test_|pccl|allreduce_|outer_step_|simplehash_kernel|crc32c_kernel)", case):
            continue  # python tests, API names and kernel names are checked elsewhere or by the build
        if case.endswith("_") or case in ("first_step",):
            continue
        # a lower-case identifier that is neither a case nor an allowed name: only flag if it looks like a test case
        if any(case.startswith(p) for p in ("wire_", "atsp_", "quantised_", "peer_", "master_", "server_", "queued_", "groups_", "strategy_")):
            missing.append(f"case:{case}")
    # pytest functions: `py:file.py::name` and `gpu:file.py::name`
    for fname, func in re.findall(r"`(?:py|gpu):(test_\w+\.py)::(\w+)", text):
        body = (REPO / "tests" / fname).read_text() if (REPO / "tests" / fname).exists() else ""
        if f"def {func}" not in body:
            missing.append(f"{fname}::{func}")
    for fname in re.findall(r"`(?:py|gpu):(test_\w+\.py)`", text):
        if not (REPO / "tests" / fname).exists():
            missing.append(fname)
    assert missing == [], missing

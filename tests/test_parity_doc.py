"""PARITY.md is the checklist against the reference's inventory: what it names must exist.

Every back-ticked source path must be a file of this repository and every back-ticked ``test_*`` name must be a test
function in tests/ (so a renamed file or test cannot silently leave the table pointing at nothing).
"""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
TEXT = (ROOT / "PARITY.md").read_text()
TOKENS = set(re.findall(r"`([^`]+)`", TEXT))


def _all_files():
    skip = (".git", "gpurun_out", "build", "__pycache__")
    return [p for p in ROOT.rglob("*") if p.is_file() and not any(s in p.parts for s in skip)]


def test_named_source_files_exist():
    files = _all_files()
    names = {p.name for p in files} | {str(p.relative_to(ROOT)) for p in files}
    missing = []
    for tok in TOKENS:
        m = re.match(r"([\w/.\-]+\.(?:cu|cuh|cpp|h|py|md|txt|patch))\b", tok)
        if not m:
            continue
        path = m.group(1)
        if path in names or any(n.endswith("/" + path) for n in names):
            continue
        missing.append(path)
    assert not missing, f"PARITY.md names files that do not exist: {sorted(missing)}"


def test_named_tests_exist():
    defined = set()
    for p in (ROOT / "tests").glob("test_*.py"):
        defined |= set(re.findall(r"^def (test_\w+)", p.read_text(), flags=re.M))
    wanted = set()
    for tok in TOKENS:
        wanted |= set(re.findall(r"\b(test_[a-zA-Z0-9_]+)\b", tok))
    wanted = {w for w in wanted if not (ROOT / "tests" / (w + ".py")).exists()}  # file names are checked above
    missing = sorted(w for w in wanted if w not in defined)
    assert not missing, f"PARITY.md names tests that do not exist: {missing}"

"""Documentation that is cheap to keep honest: every experiment switch the code reads is listed in
tools/README.md, every file under profiles/ belongs to a run tag that profiles/README.md explains."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _read(*parts):
    with open(os.path.join(ROOT, *parts), encoding="utf-8") as f:
        return f.read()


def test_every_environment_switch_is_documented():
    used = set()
    for base, _, files in os.walk(os.path.join(ROOT, "gcn_b200")):
        for fn in files:
            if fn.endswith((".cu", ".cuh", ".cpp", ".h", ".py")):
                used |= set(re.findall(r'"(GCNB_[A-Z0-9_]+)"', _read(base, fn)))
    used = {u for u in used if not u.startswith(("GCNB_ERR", "GCNB_EPI", "GCNB_OK", "GCNB_PLAN", "GCNB_F32", "GCNB_BF16"))}
    doc = _read("tools", "README.md")
    missing = sorted(u for u in used if u not in doc)
    assert not missing, f"undocumented switches: {missing}"


def test_every_profile_tag_is_indexed():
    doc = _read("profiles", "README.md")
    rows = [line.split("|")[1] for line in doc.splitlines() if line.startswith("| ") and "|" in line[2:]]
    tags = set()
    for cell in rows:
        tags |= {t.strip() for t in cell.split(",")}
    seen = {m.group(1) or m.group(2) for fn in os.listdir(os.path.join(ROOT, "profiles"))
            if (m := re.match(r"(?:r01_([a-z]+)|(r02_[a-z]+))_", fn))}
    missing = sorted(seen - tags)
    assert not missing, f"profiles/ run tags without a line in profiles/README.md: {missing}"

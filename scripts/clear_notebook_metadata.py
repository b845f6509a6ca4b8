#!/usr/bin/env python
"""Strip outputs, execution counts and volatile metadata from every ``*.ipynb`` under a directory (default: the repo), so that
notebooks diff cleanly.  Counterpart of the reference's lab/clear-metadata-notebooks.py (which drives nbconvert
preprocessors); this one works on the notebook JSON directly and needs no extra package.

  python scripts/clear_notebook_metadata.py [root] [--check]      # --check: exit 1 if a notebook would change"""
import argparse
import json
import pathlib
import sys

KEEP_NB_META = ("kernelspec", "language_info")


def clean(nb: dict) -> dict:
    for cell in nb.get("cells", []):
        cell["metadata"] = {}
        if cell.get("cell_type") == "code":
            cell["outputs"] = []
            cell["execution_count"] = None
    meta = nb.get("metadata", {})
    nb["metadata"] = {k: meta[k] for k in KEEP_NB_META if k in meta}
    if "language_info" in nb["metadata"]:
        nb["metadata"]["language_info"] = {"name": nb["metadata"]["language_info"].get("name", "python")}
    return nb


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("root", nargs="?", default=str(pathlib.Path(__file__).resolve().parent.parent))
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args(argv)
    changed = 0
    for path in sorted(pathlib.Path(a.root).rglob("*.ipynb")):
        if ".ipynb_checkpoints" in path.parts:
            continue
        old = path.read_text()
        new = json.dumps(clean(json.loads(old)), indent=1, ensure_ascii=False) + "\n"
        if new != old:
            changed += 1
            print(("would clean " if a.check else "cleaned ") + str(path))
            if not a.check:
                path.write_text(new)
    return 1 if (a.check and changed) else 0


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python
"""Fingerprints of the shipped kernels' SASS (instruction text only: no addresses, encodings or
line info), per kernel.  `--write FILE` records them, `--check FILE` lists kernels whose code
differs from the record (and new / missing ones).  Use: a number in profiles/ belongs to a
recorded build; after an edit that should not touch a measured kernel (e.g. adding an opt-in
variant next to it) `--check` proves it.

    python tools/sass_fingerprint.py --check profiles/r1_final_sass_fingerprints.json
"""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "multimodal-2025_b200", "lib", "obj")


def fingerprints():
    out = {}
    for f in sorted(os.listdir(OBJ)):
        if not f.endswith(".o"):
            continue
        sass = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, f)],
                              stdout=subprocess.PIPE, text=True, check=True).stdout
        cur = None
        for line in sass.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                cur = f + "::" + m.group(1)
                out[cur] = hashlib.sha256()
                continue
            if cur is None or not re.match(r"^\s+/\*[0-9a-f]{4,5}\*/", line):
                continue
            text = re.sub(r"/\* 0x[0-9a-f]+ \*/", "", line)
            text = re.sub(r"^\s*/\*[0-9a-f]+\*/\s*", "", text).rstrip()
            out[cur].update(text.encode() + b"\n")
    return {k: v.hexdigest()[:16] for k, v in out.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--write")
    ap.add_argument("--check")
    a = ap.parse_args()
    fp = fingerprints()
    if a.write:
        json.dump(fp, open(a.write, "w"), indent=1, sort_keys=True)
        print("recorded", len(fp), "kernels")
    if a.check:
        ref = json.load(open(a.check))
        changed = sorted(k for k in ref if k in fp and fp[k] != ref[k])
        missing = sorted(k for k in ref if k not in fp)
        new = sorted(k for k in fp if k not in ref)
        for tag, lst in (("CHANGED", changed), ("MISSING", missing), ("NEW", new)):
            for k in lst:
                print(tag, k)
        print("%d unchanged, %d changed, %d missing, %d new" %
              (len(ref) - len(changed) - len(missing), len(changed), len(missing), len(new)))
        sys.exit(1 if changed or missing else 0)


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""8-token shingle overlap of a source file with reference files: |shingles(file) & shingles(refs)| / |shingles(file)|.
Comments are stripped.  Usage: shingle_overlap.py file ref1 [ref2 ...]"""
import re, sys
def toks(path):
    s = open(path, errors="ignore").read()
    s = re.sub(r"/\*.*?\*/", " ", s, flags=re.S); s = re.sub(r"//[^\n]*", " ", s)
    return re.findall(r"[A-Za-z_]\w*|\d+\.?\d*[fF]?|\S", s)
def sh(t, k=8):
    return {tuple(t[i:i + k]) for i in range(len(t) - k + 1)}
a = sh(toks(sys.argv[1]))
for r in sys.argv[2:]:
    b = sh(toks(r))
    print("%-60s vs %-50s %5.1f %%  (%d of %d shingles)" % (sys.argv[1][-60:], r[-50:], 100.0 * len(a & b) / max(1, len(a)), len(a & b), len(a)))

#!/usr/bin/env python
"""The SASS evidence file under profiles/: every TMA / mbarrier / dependent-launch / atomic / collective / shared-store / global line of
one kernel instantiation of the in-tree libswarm_b200.so, then its opcode census.
Usage: sass_listing.py [mangled-name-substring] > profiles/rN_sass_step_warp_lean.txt"""
import collections, os, re, subprocess, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(REPO, "marl-drones_b200", "swarm_b200", "libswarm_b200.so")
want = sys.argv[1] if len(sys.argv) > 1 else "step_warp_kernelILi16EfLi1ELi5ELb1ELb1EE"
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
body = next(f for f in funcs if f.split("\n", 1)[0].find(want) >= 0)
name = body.split("\n", 1)[0].strip()
keep = re.compile(r"UBLKCP|SYNCS|UTMA|ACQBULK|PREEXIT|ATOM|RED|MATCH|SHFL|VOTE|FENCE|MEMBAR|STS|LDS\.128|STG|LDG|BAR\.|ERRBAR|CCTL|LDGSTS")
ops = collections.Counter()
lines = []
for ln in body.split("\n"):
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if not m:
        continue
    ops[m.group(1).split(".")[0]] += 1
    if keep.search(m.group(1)):
        lines.append(ln.rstrip())
print(f"# cuobjdump -sass of {name} (sm_100a) from the in-tree libswarm_b200.so: every TMA / mbarrier / dependent-launch / atomic /")
print("# collective / shared-store / global line, then the opcode census.  UBLKCP.S.G = cp.async.bulk global->shared (stage loads),")
print("# UBLKCP.G.S = the obs bulk store, SYNCS.* = mbarrier, ACQBULK = griddepcontrol.wait, PREEXIT = griddepcontrol.launch_dependents,")
print("# ATOMS.AND = the clear that elects a winner, ATOMS.ADD = the stage-release counter, REDUX.SUM = tile totals / statistics.")
print("# No tcgen05 / TMEM is expected: no stage of this path is a dense contraction.\n")
print("\n".join(lines))
print(f"\n# opcode census ({sum(ops.values())} SASS instructions in the function)")
for k, v in ops.most_common():
    print(f"{k:<12} {v}")

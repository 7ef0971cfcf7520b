"""Known answers for the muscle-signal hook, produced by THE REFERENCE'S OWN main_sim.py.

/root/reference/src/main_sim.py is Python 2; it runs under Python 3 after two mechanical rewrites applied to an
in-memory copy of its text (never written to the repo): `j = n/2` -> `j = n//2` (integer division) and
`map(` -> `list(map(` (py2 map returned a list). Asserted counts: 1 and 4. The script is then exec'ed and
muscle_simulation().run() is called; the 96 floats of calls 0, 1, 2, 500 and 1000 are stored as hex strings
(exact doubles) in tests/golden/main_sim_kat.json.

  python tests/golden/make_main_sim_kat.py     (needs /root/reference; the JSON is committed)
"""
import json
import os
import re

SRC = "/root/reference/src/main_sim.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "main_sim_kat.json")


def reference_module():
    text = open(SRC).read()
    text, n1 = re.subn(r"j = n/2\b", "j = n//2", text)
    text, n2 = re.subn(r"\(map\((.*)\)\)\s*$", r"(list(map(\1)))", text, flags=re.M)
    text, n3 = re.subn(r"= map\((.*)\)\s*$", r"= list(map(\1))", text, flags=re.M)
    assert n1 == 1 and n2 + n3 == 4, (n1, n2, n3)
    ns = {}
    exec(compile(text, SRC, "exec"), ns)
    return ns


def main():
    ns = reference_module()
    sim = ns["muscle_simulation"]()
    calls = {}
    for t in range(1001):
        v = sim.run()
        if t in (0, 1, 2, 500, 1000):
            calls[str(t)] = [float(x).hex() for x in v]
    json.dump({"source": SRC, "rewrites": ["j = n/2 -> j = n//2", "map(...) -> list(map(...)) x4"], "calls": calls},
              open(OUT, "w"), indent=0)
    print("wrote", OUT, {k: len(v) for k, v in calls.items()})


if __name__ == "__main__":
    main()

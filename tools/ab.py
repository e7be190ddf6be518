"""A/B timing of kernel variants on config 2.  `python tools/ab.py name:DEF=1,DEF2=3 name2:...` builds each variant
(wheat_3dgs_b200/variants/, done on the CPU box beforehand with --build-only) and prints stage times of each by
running tools/stage_times.py with WHEATSPLAT_LIB pointing at it."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from wheat_3dgs_b200 import build as b  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
for spec in args:
    name, _, defs = spec.partition(":")
    lib = os.path.join(ROOT, "wheat_3dgs_b200", "variants", "lib%s.so" % name)
    if "--build-only" in sys.argv:
        print(b.build_variant(name, [d for d in defs.split(",") if d]))
        continue
    env = dict(os.environ, WHEATSPLAT_LIB=lib)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "stage_times.py"), "12"], env=env,
                       capture_output=True, text=True)
    rows = dict(l.split()[:2] for l in r.stdout.splitlines() if l.strip())
    print("%-28s fwd %s bwd %s pre_bwd %s pre %s emit %s sort %s sum %s %s" % (name, rows.get("render_fwd"), rows.get("render_bwd"),
                                                        rows.get("preprocess_bwd"), rows.get("preprocess"), rows.get("emit_keys"), rows.get("sort"), rows.get("sum"),
                                                        r.stderr[-300:] if r.returncode else ""))

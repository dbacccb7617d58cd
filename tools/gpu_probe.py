"""Quick on-box probe: time the stage calls at a few box sizes and print the library's per-stage stats."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))
import numpy as np
import vrgpu, vrsynth

def main():
    sizes = [int(a) for a in sys.argv[1:]] or [64, 128]
    ctx = vrgpu.Context(0)
    for ns in sizes:
        n = ns ** 3
        t = time.time(); d = vrsynth.make_box(n, 100 + ns); tg = time.time() - t
        parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"], pinned=True)
        ell = vrsynth.BOX / ns
        b = np.sqrt((ell * ell) * (0.2 * 0.2))
        for rep in range(3):
            t = time.time(); pf, ng = ctx.fof3d(parts, b, vrsynth.BOX, 20); dt = time.time() - t
            st = ctx.stats()
            print(json.dumps({"nside": ns, "rep": rep, "ngroups": ng, "wall_s": round(dt, 4), "gen_s": round(tg, 2),
                              **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()}}), flush=True)
if __name__ == "__main__":
    main()

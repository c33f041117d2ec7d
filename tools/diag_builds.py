import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, oracle_util as ou, raytracing_course_hw_b200 as rt
for refbvh in (True, False):
    sc = ou.load_scene(ou.room_scene(20000), width=320, height=180, samples=2, reference_bvh=refbvh)
    for depth in (1, 2, 3, 6):
        out = {}
        for name, flags in (("wf", 0), ("mk", rt.FLAG_MEGAKERNEL), ("wfkeep", rt.FLAG_KEEP_DARK_RAYS)):
            r = rt.Renderer(flags=flags); r.upload(sc); r.render(0, 2, ray_depth=depth); out[name] = r.download_accum(); st = r.stats(); r.close()
            out[name + "_rays"] = st["rays"]
        d = np.abs(out["wf"] - out["mk"]).max(axis=2)
        rel = d / (np.abs(out["wf"]).max(axis=2) + 1e-6)
        print(f"refbvh={refbvh} depth={depth}: differing px {(d > 0).sum()} of {d.size}; rel>1e-5: {(rel > 1e-5).sum()}, rel>1e-2: {(rel > 1e-2).sum()}; "
              f"wf==wfkeep {np.array_equal(out['wf'], out['wfkeep'])}; means {out['wf'].mean():.6f} {out['mk'].mean():.6f}; rays {out['wf_rays']} {out['mk_rays']} {out['wfkeep_rays']}")

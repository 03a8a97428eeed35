#!/usr/bin/env python3
"""Per-ray operation counts (instrumented kernel variant, RT_RENDER_COUNT_OPS) for the given scenes and both walks.
usage: gpu_opcounts.py SCENE[:spp[:width]] ...   -> one JSON line per (scene, traversal)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402


def main():
    ctx = rt.Context(0)
    for spec in sys.argv[1:]:
        parts = [int(x) for x in spec.split(":")]
        scene = parts[0]
        spp = parts[1] if len(parts) > 1 else 8
        width = parts[2] if len(parts) > 2 else 0
        hs = rt.HostScene(scene)
        for trav in (rt.RT_TRAVERSAL_AUTO, rt.RT_TRAVERSAL_FAST):
            ctx.upload(hs, trav)
            p = hs.params(width, spp, 50, flags=rt.RT_RENDER_COUNT_OPS | rt.RT_RENDER_COUNT_RAYS)
            _, _, st = ctx.render(p)
            ops = dict(zip(rt.OP_NAMES, [int(x) for x in st.ops]))
            rays = max(1, st.rays)
            print(json.dumps({"scene": scene, "traversal": trav, "used": int(st.traversal_used), "spp": spp,
                              "paths": st.paths, "rays": st.rays, "ms": round(st.render_ms, 3),
                              "per_ray": {k: round(v / rays, 3) for k, v in ops.items()}}), flush=True)
        hs.close()
    ctx.close()


if __name__ == "__main__":
    main()

"""C5 (600 scripted moves on scene2.json, every frame frame_number = 0) through the pipelined loop: trt_frame_submit /
trt_frame_collect against trt_render_frame, and the same at 1920x1080 truecolor where read-back + a write to a file
are large enough to matter.  Wall clock of the whole loop; streams compared byte for byte.
usage: python tools/bench_c5_pipeline.py [frames]"""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 600
out = {}
for tag, (w, h, mode, n) in {"c5_200x100_ascii": (200, 100, pkg.MODE_ASCII, frames), "1920x1080_truecolor": (1920, 1080, pkg.MODE_TRUECOLOR, 60)}.items():
    scene = pkg.Scene.parse(json.dumps(dict(O.SCENE2, width=w, height=h)))
    cam = pkg.Camera()
    us = []
    for i in range(n):
        moved = cam.key(O.C5_KEYS[i % 8])
        us.append(scene.uniforms(cam, O.c2_seed(i), 0 if moved else 1))
    res = {}
    streams = {}
    for kind in ("serial", "pipelined"):
        r = pkg.Renderer(w, h, conv=2)
        r.upload_scene(scene)
        sink = open(os.path.join(tempfile.gettempdir(), f"trt_{tag}_{kind}.bin"), "wb")
        for rep in range(2):                                  # first repetition warms up (NVRTC, buffers)
            sink.seek(0); sink.truncate()
            t0 = time.perf_counter()
            if kind == "serial":
                for u in us:
                    sink.write(r.render_frame(u, mode))
            else:
                r.frame_submit(us[0], mode)
                for u in us[1:]:
                    r.frame_submit(u, mode)
                    sink.write(r.frame_collect())
                sink.write(r.frame_collect())
            sink.flush()
            dt = time.perf_counter() - t0
        sink.close()
        streams[kind] = open(sink.name, "rb").read()
        res[kind] = {"ms_total": dt * 1e3, "ms_per_frame": dt * 1e3 / n, "fps": n / dt, "bytes_per_frame": len(streams[kind]) // n}
        r.close()
    res["identical_streams"] = streams["serial"] == streams["pipelined"]
    res["speedup"] = res["serial"]["ms_total"] / res["pipelined"]["ms_total"]
    out[tag] = res
print(json.dumps(out))

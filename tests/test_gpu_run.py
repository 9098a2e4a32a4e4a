"""GPU tests of the reference's public entry `run(full_color, verbose, scene_path)` (lib.rs:100) as restated in
trt_run.cpp: the frame loop (lib.rs:386-566) headless into a file, with scripted keys (lib.rs:392-412), with
the frame-batch extension, and on a real pseudo-terminal through the CLI binary (raw mode, key polling, Esc).
Bar: the byte stream written for every displayed frame equals the oracle's frame loop bit for bit (strict
convention), framing as lib.rs:497,532."""
import json
import os
import select
import subprocess
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
HOME = b"\x1b[1;1H"
GOLDEN = 0x9E3779B9


def _expected(O, sc, w, h, seed, keys, frames, mode, shown=None):
    """The oracle's loop: one key per frame, reset on a move, seed_f = seed + GOLDEN * f."""
    prims = O.flatten_scene(sc)
    cam = O.Camera()
    acc, frame_count, moved_flag, out = None, 0, True, []
    for f in range(frames):
        k = keys[f] if f < len(keys) else ""
        if k and cam.key(k):
            moved_flag, frame_count = True, 0
        u = O.make_uniforms(sc, len(prims), (seed + GOLDEN * f) & 0xFFFFFFFF, 0 if moved_flag else frame_count, camera=cam,
                            width=w, height=h)
        ref = O.render(prims, u, accumulation=acc)
        acc = ref["accumulation"]
        frame_count += 1
        moved_flag = False
        if shown is None or f in shown:
            out.append(HOME + O.quantise(ref["pixels"], w, h, mode)[0])
    return b"".join(out)


@pytest.mark.parametrize("full_color", [False, True])
def test_run_headless_matches_oracle_loop(pkg, O, tmp_path, full_color):
    path = os.path.join(SCENES, "scene2.json")
    sc = json.load(open(path))
    out = str(tmp_path / "frames.bin")
    rc = pkg.run(full_color, False, path, width=60, height=30, frames=4, seed=5, out_path=out, quiet=1, conv=0)
    assert rc == 0
    assert open(out, "rb").read() == _expected(O, sc, 60, 30, 5, [], 4, 1 if full_color else 0)


def test_run_scripted_keys_reset_accumulation(pkg, O, tmp_path):
    path = os.path.join(SCENES, "scene2.json")
    sc = json.load(open(path))
    out = str(tmp_path / "frames.bin")
    keys = ["w", "Right", "d", "Up"]
    rc = pkg.run(False, False, path, width=48, height=24, frames=6, seed=9, keys=",".join(keys), out_path=out, quiet=1, conv=0)
    assert rc == 0
    assert open(out, "rb").read() == _expected(O, sc, 48, 24, 9, keys, 6, 0)


def test_run_batched_frames_show_the_last_of_each_batch(pkg, O, tmp_path):
    path = os.path.join(SCENES, "Cornell_Box.json")
    sc = dict(json.load(open(path)), samples_per_pixel=16, max_depth=6)
    spath = str(tmp_path / "scene.json")
    json.dump(sc, open(spath, "w"))
    out = str(tmp_path / "frames.bin")
    rc = pkg.run(True, False, spath, width=40, height=20, frames=6, seed=3, out_path=out, quiet=1, conv=0, batch_frames=3)
    assert rc == 0
    assert open(out, "rb").read() == _expected(O, sc, 40, 20, 3, [], 6, 1, shown={2, 5})


def test_run_status_line_format(pkg, tmp_path):
    path = os.path.join(SCENES, "scene2.json")
    out = str(tmp_path / "frames.bin")
    assert pkg.run(False, False, path, width=32, height=16, frames=2, seed=1, out_path=out, conv=2) == 0
    data = open(out, "rb").read()
    # lib.rs:551-555: cursor to row height + 1, then the fields of the reference's status line
    assert data.count(b"\x1b[17;1HFrame: ") == 2
    for field in (b"Frame: 2/", b" | FPS: ", b" | GPU: ", b"ms | CPU: 0ms | Cleanup: 0ms | Other: ", b"ms | Total: "):
        assert field in data


def test_cli_on_a_pseudo_terminal(pkg):
    """The drop-in binary on a real (pseudo) terminal: frame clamped to the terminal size (lib.rs:114-115), raw
    mode and hidden cursor (lib.rs:355-368), keys polled between frames, Esc leaves the loop, the cursor is
    shown again and "Exiting." is printed (lib.rs:566-570)."""
    import fcntl
    import pty
    import struct
    import termios
    exe = os.path.join(ROOT, "terminal-raytracer_b200", "terminal-raytracer")
    assert os.path.exists(exe), "CLI binary missing (make -C terminal-raytracer_b200/csrc)"
    master, slave = pty.openpty()
    fcntl.ioctl(slave, termios.TIOCSWINSZ, struct.pack("HHHH", 24, 80, 0, 0))          # 24 rows x 80 columns
    p = subprocess.Popen([exe, "--path", os.path.join(SCENES, "scene2.json"), "--seed", "1"], stdin=slave, stdout=slave,
                         stderr=subprocess.PIPE, close_fds=True)
    os.close(slave)
    chunks = []

    def pump(seconds):
        end = time.time() + seconds
        while time.time() < end:
            r, _, _ = select.select([master], [], [], 0.05)
            if r:
                try:
                    chunk = os.read(master, 1 << 20)
                except OSError:
                    return
                if not chunk:
                    return
                chunks.append(chunk)

    err = b""
    try:
        # the first frame waits for the CUDA context, the module load and the run-time compile of the scene's kernel: seconds
        # on a fresh box, so poll for it instead of assuming a start-up time
        n_before = 0
        for _ in range(120):
            pump(0.5)
            n_before = b"".join(chunks).count(b"Frame: ")
            if n_before >= 1 or p.poll() is not None:
                break
        assert n_before >= 1, b"".join(chunks)[:200]
        os.write(master, b"w")                     # move: the accumulation restarts at frame 1
        pump(0.5)
        os.write(master, b"\x1b")                  # Esc
        for _ in range(100):                       # drain until the slave side closes (EIO) or the process ends
            pump(0.1)
            if p.poll() is not None:
                break
            time.sleep(0.05)
        assert p.wait(timeout=30) == 0, "the binary did not leave its loop on Esc"
    finally:
        if p.poll() is None:
            p.kill()
        err = p.stderr.read()
        os.close(master)
    buf = b"".join(chunks)
    assert b"\x1b[?25l" in buf[:64] and b"\x1b[2J" in buf[:64]      # cursor hidden, screen cleared
    assert b"\x1b[?25h" in buf[-4096:]                               # cursor shown again
    assert b"Exiting." in err
    # scene2 is 200x100: clamped to 80 columns x 22 rows -> the status line sits on row 23
    assert b"\x1b[23;1HFrame: " in buf
    assert buf.count(b"Frame: 1/") >= 2                              # the first frame, and the first one after the move

"""TEST INFRASTRUCTURE ONLY (oracle).  Python driver of oracle/_ref/ref_raybatch --serve, the
UNMODIFIED reference kd-tree compiled from /root/reference by oracle/Makefile.  Imported only by
tests/, __graft_entry__.smoke() and bench.py's CPU legs (cpu_baseline / --impl reference); never by
the product package."""
import json
import os
import subprocess
import tempfile

import numpy as np

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
HIT_DTYPE = np.dtype([("t", "f4"), ("bu", "f4"), ("bv", "f4"), ("prim", "i4")])


def have_ref(name="ref_raybatch"):
    return os.path.exists(os.path.join(REF_DIR, name))


class RefServer:
    """Long-lived reference kd-tree (oracle/_ref/ref_raybatch --serve): builds the tree once,
    then traces ray batches on the host cores.  Used by bench.py's CPU legs and the larger
    parity tests so that the serial kd build is paid once."""

    def __init__(self, tris, threads=0, accel="yafaray-kdtree-original"):
        self._dir = tempfile.TemporaryDirectory()
        d = self._dir.name
        tp = os.path.join(d, "t.f32")
        np.ascontiguousarray(tris, np.float32).tofile(tp)
        cmd = [os.path.join(REF_DIR, "ref_raybatch"), tp, "--serve", "-", "-", str(int(threads)), accel]
        self._p = subprocess.Popen(cmd, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, bufsize=1)
        self.info = json.loads(self._p.stdout.readline())
        self._k = 0

    def trace(self, rays, kind, repeats=1):
        d = self._dir.name
        self._k += 1
        rp, hp = os.path.join(d, f"r{self._k}.f32"), os.path.join(d, f"h{self._k}.bin")
        np.ascontiguousarray(rays, np.float32).tofile(rp)
        self._p.stdin.write(f"{kind} {rp} {hp} {int(repeats)}\n")
        self._p.stdin.flush()
        line = self._p.stdout.readline()
        if not line:
            raise RuntimeError("ref_raybatch --serve died")
        hits = np.fromfile(hp, HIT_DTYPE)
        os.remove(rp)
        os.remove(hp)
        return hits, json.loads(line)

    def close(self):
        if self._p and self._p.poll() is None:
            try:
                self._p.stdin.close()
                self._p.wait(timeout=30)
            except Exception:
                self._p.kill()
        self._p = None
        self._dir.cleanup()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class ScriptRecorder:
    """Records C-API style calls (method name == C function minus `yafaray_`) as ref_render script lines."""

    def __init__(self):
        self.lines = []

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)

        def call(*args):
            toks = [name]
            for a in args:
                if isinstance(a, bool):
                    toks.append("1" if a else "0")
                elif isinstance(a, float):
                    toks.append(repr(float(a)))
                else:
                    toks.append(str(a))
            self.lines.append(" ".join(toks))

        return call


def ref_render(recorder, env=None, cwd=None):
    """Render a recorded scene with the UNMODIFIED reference -> (float32 [H, W, 4] linear RGBA, info).
    cwd: working directory of the renderer (relative image file names of textures resolve against it)."""
    with tempfile.TemporaryDirectory() as d:
        sp, op = os.path.join(d, "scene.script"), os.path.join(d, "out.f32")
        with open(sp, "w") as f:
            f.write("\n".join(recorder.lines) + f"\nrender {op}\n")
        e = dict(os.environ)
        e.update(env or {})
        out = subprocess.run([os.path.join(REF_DIR, "ref_render"), sp], capture_output=True, text=True, check=True, env=e, cwd=cwd)
        info = json.loads(out.stdout.strip().splitlines()[-1])
        img = np.fromfile(op, np.float32).reshape(info["height"], info["width"], 4)
        return img, info

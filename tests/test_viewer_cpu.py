"""CPU: the headless viewer's camera (tools/gltf_viewer/FlyCamera.hpp) reproduces the camera BEHAVIOUR of the reference's viewer
(src/gltf_viewer/CameraController.hpp:100-173: look with the right button held, zoom by scroll, velocity eased towards the held
direction and back to rest, motion in the camera's frame) under scripted input.  The expected poses come from an independent
restatement below, written the way the reference does it -- axis-angle rotations of the forward vector -- in float64; the C++
works in float32 with cross products, so poses agree to rounding, not bit for bit."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
CPP = ROOT / "tests" / "cpp"
HALF_PI = np.pi / 2


@pytest.fixture(scope="module")
def tool():
    subprocess.run(["make", "-C", str(CPP), "cpu"], check=True, capture_output=True)
    return str(CPP / "build" / "fly_camera")


def rotate(v, axis, angle):  # Rodrigues
    k = axis / np.linalg.norm(axis)
    return v * np.cos(angle) + np.cross(k, v) * np.sin(angle) + k * np.dot(k, v) * (1 - np.cos(angle))


class Model:
    def __init__(self):
        self.p, self.v = np.zeros(3), np.zeros(3)
        self.yaw, self.pitch, self.fov = -HALF_PI, 0.0, HALF_PI
        self.cursor = np.zeros(2)
        self.keys = (0, 0, 0, False)

    def direction(self):
        c = np.cos(self.pitch)
        d = np.array([c * np.cos(self.yaw), np.sin(self.pitch), c * np.sin(self.yaw)])
        return d / np.linalg.norm(d)

    def step(self, op, a):
        if op == "place":
            self.p = np.array(a[:3], float)
            self.yaw, self.pitch, self.fov = a[3], float(np.clip(a[4], -HALF_PI + 0.01, HALF_PI - 0.01)), float(np.clip(a[5], 0.1, HALF_PI))
        elif op == "keys":
            self.keys = (int(a[0]), int(a[1]), int(a[2]), a[3] != 0)
        elif op == "cursor":
            new = np.array(a, float)
            delta, self.cursor = new - self.cursor, new
            if self.keys[3]:
                self.yaw -= delta[0] * 0.01
                self.pitch = float(np.clip(self.pitch - delta[1] * 0.01, -HALF_PI + 0.01, HALF_PI - 0.01))
        elif op == "scroll":
            self.fov = float(np.clip(self.fov - a[0] * 0.1, 0.1, HALF_PI))
        elif op == "advance":
            for _ in range(int(a[1]) if len(a) > 1 else 1):
                lerp = a[0] * 10.0
                wish = np.array(self.keys[:3], float)
                if np.linalg.norm(wish) != 0:
                    self.v = self.v + (wish / np.linalg.norm(wish) * 0.02 - self.v) * lerp
                fwd = self.direction()
                right = np.cross(fwd, [0.0, 1.0, 0.0])
                up = rotate(fwd, right, HALF_PI)
                col0 = rotate(fwd, up, HALF_PI)
                self.p = self.p + col0 * self.v[0] + up * self.v[1] + fwd * self.v[2]
                self.v = self.v + (0 - self.v) * lerp

    def pose(self):
        return np.array([*self.p, self.yaw, self.pitch, self.fov, *self.v])


SCRIPTS = [
    ["keys:0,0,1,0", "advance:0.016,30", "keys:0,0,0,0", "advance:0.016,20"],                                   # W, then coast to rest
    ["keys:-1,1,0,1", "cursor:40,-25", "advance:0.02,10", "cursor:140,80", "advance:0.02,10", "scroll:3", "scroll:-20"],
    ["place:1,-2,3,0.7,0.4,1.0", "keys:1,-1,-1,1", "cursor:-300,900", "advance:0.01,25", "cursor:-300,-2000", "advance:0.05,4", "scroll:100"],
    ["keys:1,0,1,0", "advance:0.25,3"],                                                                          # a long frame: the blend overshoots
    ["cursor:500,500", "keys:0,0,0,1", "cursor:510,495", "advance:0.016,1"],                                     # no look before the button is held
]


@pytest.mark.parametrize("script", SCRIPTS)
def test_camera_poses_after_scripted_input(tool, script):
    r = subprocess.run([tool, *script], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    model = Model()
    for step, line in zip(script, lines):
        op, _, args = step.partition(":")
        model.step(op, [float(t) for t in args.split(",")])
        got = np.array([float(t) for t in line.split()])
        assert np.allclose(got, model.pose(), rtol=2e-5, atol=2e-6), f"after {step}: {got} vs {model.pose()}"
    # the view matrix looks along the modelled direction from the modelled position (makeCameraData: RH, up = -Y)
    view = np.array([float(t) for t in lines[-1].split()[1:]]).reshape(4, 4).T  # printed column-major
    fwd = model.direction()
    assert np.allclose(-view[2, :3], fwd, atol=1e-5)
    assert np.allclose(view[:3, :3] @ model.p + view[:3, 3], 0, atol=1e-4)

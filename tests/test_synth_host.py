"""Host-side properties of the synthetic scenes (scorp_b200/synth.py): runs without a GPU."""
import numpy as np
import torch

from scorp_b200 import synth


def test_scene_generation_is_independent_of_the_thread_count():
    """make_scene must give the same bits whatever torch.get_num_threads() is (it pins one thread internally)."""
    import hashlib
    def sha(sc):
        hh = hashlib.sha256()
        for k in ("means3D", "scales", "rotations", "opacities", "shs"):
            hh.update(getattr(sc, k).numpy().tobytes())
        return hh.hexdigest()
    n0 = torch.get_num_threads()
    try:
        got = set()
        for n in (1, 2, 5, 8):
            torch.set_num_threads(n)
            got.add(sha(synth.make_scene("C2", P=150_001)))
            assert torch.get_num_threads() == n          # restored
    finally:
        torch.set_num_threads(n0)
    assert len(got) == 1


def test_stress_camera_sits_inside_the_cloud():
    cam = synth.stress_camera(640, 400, 3)
    assert torch.allclose(cam.campos, torch.tensor(synth.CENTER), atol=1e-5)
    sc = synth.make_scene("C2", P=20_000, W=640, H=400, cameras=[cam])
    z = (torch.cat([sc.means3D, torch.ones(sc.P, 1)], 1) @ cam.viewmatrix)[:, 2]
    assert 0.4 < float((z > 0.2).float().mean()) < 0.6          # about half of the Gaussians are behind the camera


def test_c5_sweep_uses_the_reference_rotations():
    from pathlib import Path
    rots = np.load(Path(__file__).resolve().parent / "golden" / "rotations_128.npz")["rotations"]
    assert rots.shape == (128, 3, 3) and rots.dtype == np.float64
    assert np.allclose(rots @ rots.transpose(0, 2, 1), np.eye(3), atol=1e-9)      # proper rotations
    assert synth.CONFIGS["C5"]["sweep_views"] == 128
    cam = synth.orbit_camera(512, 512, rots[5])
    c = torch.tensor(synth.CENTER + (1.0,))
    assert abs(float((c @ cam.viewmatrix)[2]) - 6.0) < 1e-5        # the scene centre stays at depth 6

"""A synthetic extrinsic-calibration scene for the pose-refinement tests: a gently curved wall whose LiDAR intensity is a smooth function
of the position on the wall, seen by `n_frames` cameras whose images render that same function.  NID is then minimal at the true
extrinsic (identity), and a perturbed initial guess has to be pulled back."""
import numpy as np

INTR = (300.0, 300.0, 160.0, 120.0)
DIST = (0.0, 0.0, 0.0, 0.0, 0.0)
W, H = 320, 240


def _texture(x, y):
    return 0.5 + 0.22 * np.sin(5.0 * x + 0.7) * np.cos(4.0 * y - 0.3) + 0.2 * np.sin(2.3 * x * y + 1.1) + 0.05 * np.cos(11.0 * x)


def make(n_points=30_000, n_frames=2, seed=9):
    """-> list of (culled cloud in the camera frame n x 4 float32, image plane H x W float64 in [0,1])."""
    rng = np.random.default_rng(seed)
    frames = []
    for f in range(n_frames):
        depth = 3.0 + 0.4 * f
        # pixel grid -> wall (a plane at z = depth: exact back-projection for the distortion-free camera)
        u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))  # NIDCost samples pixel i at coordinate i
        xw, yw = (u - INTR[2]) / INTR[0] * depth, (v - INTR[3]) / INTR[1] * depth
        img = np.clip(_texture(xw + 0.3 * f, yw), 0.0, 1.0)
        img = np.round(img * 255.0) / 255.0                      # an 8-bit image / 255, like convertTo(CV_64F, 1/255)
        px = rng.uniform(8, W - 8, n_points)
        py = rng.uniform(8, H - 8, n_points)
        x, y = (px - INTR[2]) / INTR[0] * depth, (py - INTR[3]) / INTR[1] * depth
        inten = np.clip(_texture(x + 0.3 * f, y) + rng.normal(0, 0.01, n_points), 0.0, 0.999)
        pts = np.stack([x, y, np.full(n_points, depth), inten], 1).astype(np.float32)
        frames.append((pts, img))
    return frames


def perturbation():
    """Initial T_camera_lidar: 8 mm / 0.25 degrees away from the truth (identity), well inside MultiNIDCost's gate."""
    from scipy.spatial.transform import Rotation
    T = np.eye(4)
    T[:3, :3] = Rotation.from_rotvec(np.deg2rad([0.15, -0.18, 0.08])).as_matrix()
    T[:3, 3] = [0.006, -0.005, 0.002]
    return T


def pose_error(T):
    from scipy.spatial.transform import Rotation
    return float(np.linalg.norm(T[:3, 3])), float(Rotation.from_matrix(T[:3, :3]).magnitude())

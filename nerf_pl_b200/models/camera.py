"""Host-side camera bookkeeping for the shadow kernels.

A view is described by its centre of projection `eye_pos` and a 3x3 matrix `camera` whose columns are the pixel-step
vectors a (one pixel to the right), b (one pixel down) and the vector c from the eye to the image's top-left corner,
so that the ray through pixel (u, v) has direction a*u + b*v + c.  The class keeps the public surface the shadow-mapping
trainers use (reference models/camera.py:5-132: constructor from a horizontal field of view, posing by a Blender
camera-to-world matrix or by look-at, the change of frame between two views); the arithmetic lives in the small
functions below.  Nothing here runs per ray: the kernels receive the nine numbers."""
import numpy as np
import torch

_BLENDER_TO_STANDARD = ((1, 0, 0, 0), (0, 0, -1, 0), (0, 1, 0, 0), (0, 0, 0, 1))   # x right, y in, z up -> x right, y up, z out


def _normalised(v):
    return v / torch.linalg.norm(v)


def canonical_matrix(hfov_deg, res):
    """Columns (a, b, c) of an unposed view: unit pixel steps, image plane at the distance that gives the horizontal
    field of view `hfov_deg` over `res[0]` pixels (reference :20-31)."""
    width, height = res
    half_angle = torch.tensor(hfov_deg) / torch.tensor(180.0) * np.pi / 2.0
    plane_dist = width / (2 * torch.tan(half_angle))
    a = torch.tensor([1.0, 0.0, 0.0])
    b = torch.tensor([0.0, -1.0, 0.0])
    c = torch.tensor([-width / 2.0, height / 2, -plane_dist])
    return torch.stack([a, b, c]).T


def lookat_frame(eye, target, up_hint):
    """Right / up / back unit vectors (numpy) of a camera at `eye` facing `target` (reference :50-67)."""
    back = eye - target
    back = back / np.linalg.norm(back)
    right = np.cross(up_hint, back)
    right = right / np.linalg.norm(right)
    return right, np.cross(back, right), back


def frame_change(src_M, src_eye, dst_M, dst_eye, device):
    """(R, Q): a point w * (src_M @ pix) + src_eye expressed in the destination view's (u*w', v*w', w') coordinates is
    w * (R @ pix) + Q, with R = dst_M^-1 src_M and Q = dst_M^-1 (src_eye - dst_eye) (reference :121-132)."""
    dst_inv = torch.inverse(dst_M).to(device)
    offset = (src_eye.to(device) - dst_eye.to(device)).to(device)
    return dst_inv @ src_M.to(device), dst_inv @ offset


class Camera():
    def __init__(self, hfov, res):
        self.camera = self.initialize_camera_matrix(hfov, res)
        self.res = res
        self._coord_trans = torch.tensor(_BLENDER_TO_STANDARD, dtype=torch.float32)

    def initialize_camera_matrix(self, hfov, res):
        return canonical_matrix(hfov, res)

    # columns of the matrix
    def get_a(self):
        return self.camera[:, 0]

    def get_b(self):
        return self.camera[:, 1]

    def get_c(self):
        return self.camera[:, 2]

    @classmethod
    def from_camera_eyepos(cls, eye_pos, camera):
        """A view from an already posed matrix and eye position (what a batch carries, reference :42-48)."""
        view = cls(30, (400, 400))
        view.res, view.camera, view.eye_pos = None, camera, eye_pos
        return view

    @staticmethod
    def c2w_from_lookat(eye_pos, look_at_point, up_guidance=np.array([0, 1, 0], dtype=np.float32)):
        """4x4 camera-to-world matrix (float32 numpy) of a look-at pose."""
        right, up, back = lookat_frame(eye_pos, look_at_point, up_guidance)
        c2w = np.eye(4, dtype=np.float32)
        for col, vec in enumerate((right, up, back, eye_pos)):
            c2w[:3, col] = vec
        return c2w

    def set_pose_using_blender_matrix(self, c2w, transform_coords=False):
        """Pose by a (3,4) camera-to-world matrix: the eye is its last column, the matrix is rotated by its first three
        (reference :69-89; `transform_coords` is retired there and refused here alike)."""
        if transform_coords:
            raise ValueError("This is not needed anymore. please do not use this flag.")
        rotation, translation = c2w[:, :3].float(), c2w[:, 3].float()
        self.eye_pos = translation
        self.camera = rotation @ self.camera.float()

    def set_camera_matrix(self, eye_pos, lookAtPoint, upGuidance):
        """Pose by eye / target / up hint, keeping the focal length of the current matrix (reference :91-119)."""
        width, height = self.res
        self.upGuidance, self.lookAtPoint, self.eye_pos = (torch.tensor(upGuidance), torch.tensor(lookAtPoint),
                                                           torch.tensor(eye_pos))
        view_dir = _normalised(self.lookAtPoint - self.eye_pos)
        a = _normalised(torch.cross(view_dir, self.upGuidance, dim=-1))
        b = _normalised(torch.cross(view_dir, a, dim=-1))
        plane_normal = _normalised(torch.cross(self.get_a(), self.get_b(), dim=-1))
        focal = torch.dot(plane_normal, self.get_c())
        c = view_dir * focal - a * width / 2.0 - b * height / 2.0
        self.camera = torch.stack([a, b, c]).T

    def get_transformation_to(self, to_camera, device='cpu'):
        return frame_change(self.camera, self.eye_pos, to_camera.camera, to_camera.eye_pos, device)

"""Planar-pinhole camera (PPC) algebra with the interface of models/camera.py:5-132.

Host-side 3x3 bookkeeping that feeds the shadow kernels; nothing here is on the hot path."""
import numpy as np
import torch


class Camera():
    def __init__(self, hfov, res):
        self.camera = self.initialize_camera_matrix(hfov, res)
        self.res = res
        # Blender (x right, y in, z up) -> standard (x right, y up, z out); reference :18
        self._coord_trans = torch.tensor([[1, 0, 0, 0], [0, 0, -1, 0], [0, 1, 0, 0], [0, 0, 0, 1]],
                                         dtype=torch.float32)

    def initialize_camera_matrix(self, hfov, res):
        """M = [a b c] (columns), reference :20-31."""
        w, h = res
        half = torch.tensor(hfov) / torch.tensor(180.0) * np.pi / 2.0
        cols = [torch.tensor([1.0, 0.0, 0.0]), torch.tensor([0.0, -1.0, 0.0]),
                torch.tensor([-w / 2.0, h / 2, -w / (2 * torch.tan(half))])]
        return torch.stack(cols).T

    def get_a(self):
        return self.camera[:, 0]

    def get_b(self):
        return self.camera[:, 1]

    def get_c(self):
        return self.camera[:, 2]

    @classmethod
    def from_camera_eyepos(cls, eye_pos, camera):
        """reference :42-48"""
        obj = cls(30, (400, 400))
        obj.res = None
        obj.camera = camera
        obj.eye_pos = eye_pos
        return obj

    @staticmethod
    def c2w_from_lookat(eye_pos, look_at_point, up_guidance=np.array([0, 1, 0], dtype=np.float32)):
        """4x4 camera-to-world for a camera at eye_pos looking at look_at_point (reference :50-67)."""
        back = eye_pos - look_at_point
        back = back / np.linalg.norm(back)
        right = np.cross(up_guidance, back)
        right = right / np.linalg.norm(right)
        up = np.cross(back, right)
        m = np.eye(4, dtype=np.float32)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, up, back, eye_pos
        return m

    def set_pose_using_blender_matrix(self, c2w, transform_coords=False):
        """c2w (3,4): eye = c2w[:,3], M <- c2w[:, :3] @ M (reference :69-89)."""
        if transform_coords:
            raise ValueError("This is not needed anymore. please do not use this flag.")
        self.eye_pos = c2w[:, 3].float()
        self.camera = c2w[:, :3].float() @ self.camera.float()

    def set_camera_matrix(self, eye_pos, lookAtPoint, upGuidance):
        """reference :91-119"""
        w, h = self.res
        self.upGuidance = torch.tensor(upGuidance)
        self.lookAtPoint = torch.tensor(lookAtPoint)
        self.eye_pos = torch.tensor(eye_pos)
        unit = lambda v: v / torch.linalg.norm(v)
        vd = unit(self.lookAtPoint - self.eye_pos)
        a = unit(torch.cross(vd, self.upGuidance, dim=-1))
        b = unit(torch.cross(vd, a, dim=-1))
        focal = torch.dot(unit(torch.cross(self.camera[:, 0], self.camera[:, 1], dim=-1)), self.camera[:, 2])
        c = vd * focal - a * w / 2.0 - b * h / 2.0
        self.camera = torch.stack([a, b, c]).T

    def get_transformation_to(self, to_camera, device='cpu'):
        """(R, Q) with R = ML^-1 MC, Q = ML^-1 (OC - OL) (reference :121-132)."""
        inv = torch.inverse(to_camera.camera).to(device)
        delta = (self.eye_pos.to(device) - to_camera.eye_pos.to(device)).to(device)
        return inv @ self.camera.to(device), inv @ delta

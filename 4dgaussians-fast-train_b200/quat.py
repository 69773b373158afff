"""Quaternion product used by the `apply_rotation` option (reference: utils/graphics_utils.py:109-131)."""
import torch


def batch_quaternion_multiply(q1, q2):
    """Hamilton product of two [N,4] (w,x,y,z) batches, normalised."""
    w1, x1, y1, z1 = q1.unbind(dim=1)
    w2, x2, y2, z2 = q2.unbind(dim=1)
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    q3 = torch.stack((w, x, y, z), dim=1)
    return q3 / torch.norm(q3, dim=1, keepdim=True)

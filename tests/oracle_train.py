"""Trainer::train's view loop (reference src/Trainer.cu:311-420) on the CPU oracle -- shared by tests."""
import numpy as np

from oracle import pyoracle


def oracle_running_means(c, cloud, cams, truthW, truthB, samples=None, M=4):
    """Running means [var, loc, sh, scale, opacity, rot] over the white then black passes of `cams`.
    `samples` is S of accumulateGradients (2 x the camera count over ALL ranks); default 2 * len(cams)."""
    P, W, H = c["P"], c["W"], c["H"]
    n = len(cams)
    S = float(samples if samples is not None else 2 * n)
    avg = [np.zeros(k, np.float32) for k in (P, 3 * P, 3 * M * P, 3 * P, P, 4 * P)]
    for i in range(2 * n):
        white = i < n
        cs = cams[i % n].view_constants(W, H, square_fov=(W == H))
        bg = (1.0, 1.0, 1.0) if white else (0.0, 0.0, 0.0)
        out, v = pyoracle.forward(P, 1, M, bg, W, H, cloud["locations"], cloud["shs"], None, cloud["opacities"], cloud["scales"],
                                  1.0, cloud["rotations"], None, cs["view"], cs["projview"], cs["location"], cs["tan_fovx"],
                                  cs["tan_fovy"])
        truth = np.ascontiguousarray((truthW if white else truthB)[i % n]).view(np.uint32)
        loss = pyoracle.image_int_to_loss(truth, out, W, H)
        g = v.backward(loss, M)
        pyoracle.accumulate(avg, [g["dL_dmean3D"], g["dL_dsh"], g["dL_dscale"], g["dL_dopacity"], g["dL_drot"]], S, M, P)
        v.close()
    return avg

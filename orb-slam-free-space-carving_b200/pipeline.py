"""Batch driver of the semi-dense path over a key-frame sequence: the loop structure of
ProbabilityMapping::SemiDenseLoop (PM.cc:137-318) on top of the C ABI.

  phase 1  per key frame: upload + gradient front end (K0)
  phase 2  per key frame: neighbour pick, in-plane rotation, StereoSearchConstraints (host),
           then epipolar search + fusion + intra-KF checks (K1..K3)          PM.cc:151-244
  phase 3  per key frame whose neighbours all have maps: inter-KF check (K4) and point set (K5)
                                                                              PM.cc:262-307
"""
import numpy as np

from . import Context, stereo_search_constraints
from .synth import Scene


class SequencePlan:
    """Host-side inputs of a synthetic sequence (what ORB-SLAM's map would provide)."""

    def __init__(self, scene: Scene, nkf: int, n_nbr: int, first_kf: int = 0, sparse_points: int = 1000,
                 prior_sigmas: float = 2.0):
        self.scene, self.nkf, self.n_nbr, self.first = scene, nkf, n_nbr, first_kf
        self.ids = list(range(first_kf, first_kf + nkf))
        self.poses = {k: scene.pose(k) for k in self.ids}
        self.nbrs, self.rots, self.prior = {}, {}, {}
        for k in self.ids:
            nb = Scene.neighbours(k - first_kf, nkf, n_nbr) + first_kf
            self.nbrs[k] = nb.astype(np.int32)
            self.rots[k] = np.array([Scene.rot_in_plane(k, j) for j in nb], np.float32)
            d = scene.sparse_depths(k, sparse_points)
            if prior_sigmas == 2.0:
                self.prior[k] = stereo_search_constraints(d)  # PM.cc:370-383 verbatim
            else:  # BASELINE config 5: widened prior, mean +- s*sigma
                mean, sd = float(np.mean(d)), float(np.std(d))
                self.prior[k] = (float(np.float32(1.0 / (mean - prior_sigmas * sd))),
                                 float(np.float32(1.0 / (mean + prior_sigmas * sd))))


def upload_all(ctx: Context, plan: SequencePlan, images):
    for k in plan.ids:
        ctx.upload_keyframe(k, images[k], plan.scene.K, plan.poses[k])


def semidense_all(ctx: Context, plan: SequencePlan):
    for k in plan.ids:
        mn, mx = plan.prior[k]
        ctx.semidense(k, plan.nbrs[k], plan.rots[k], mn, mx)


def interkf_all(ctx: Context, plan: SequencePlan, pointset=True):
    for k in plan.ids:
        ctx.interkf_check(k, plan.nbrs[k])
        if pointset:
            ctx.update_pointset(k)

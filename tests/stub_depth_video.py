"""Test stand-in for the reference's BA dispatcher `DepthVideo.ba / .dspo`.

TEST INFRASTRUCTURE.  The reference file src/depth_video.py cannot be imported on the GPU box (it
is not there, and it pulls in the network, lietorch and the dataset code), so its dispatch logic
is restated here statement by statement -- the same tensors, the same permute, the same 16
positional arguments, the same clamp and the same re-binding of the state -- with the modules it
calls (`droid_backends`, `src.geom.ba`, `align_scale_and_shift`) resolved to this repository's
drop-ins exactly as a GlORIE-SLAM checkout would resolve them once glorie-slam_b200/ is first on
sys.path.  Lines cited are src/depth_video.py of the reference.
"""
import torch

import common
import droid_backends
import geom_ba


class SE3Stub:
    """lietorch.SE3 as far as depth_video.py:221,281 uses it: wraps [1,N,7] data, .vec() returns it"""
    manifold_dim = 6

    def __init__(self, data):
        self.data = data

    def vec(self):
        return self.data


class StubDepthVideo:
    def __init__(self, poses, disps, intrinsics, mono_disps, n_keyframes, mono_thres=0.1, visible_num=2, mv_thresh=0.05,
                 ba_type="DSPO"):
        n, ht, wd = disps.shape  # buffers are longer than the number of keyframes in use (:30-52)
        self.device = poses.device
        self.ht, self.wd, self.down_scale = ht * 8, wd * 8, 8
        self.poses, self.disps = poses, disps
        self.intrinsics = intrinsics[None].repeat(n, 1)
        self.zeros = torch.zeros_like(disps)                                      # disps_sens (:217)
        self.mono_disps = mono_disps
        self.depth_scale = torch.ones(n, device=self.device)
        self.depth_shift = torch.zeros(n, device=self.device)
        self.valid_depth_mask_small = torch.zeros_like(disps, dtype=torch.bool)
        self.counter_value = n_keyframes
        self.mono_thres, self.visible_num, self.mv_thresh, self.BA_type = mono_thres, visible_num, mv_thresh, ba_type
        self.calls = []

    # :326-361 with up=False
    def update_valid_depth_mask(self):
        curr_idx = self.counter_value - 1
        dirty_index = torch.arange(curr_idx + 1).to(self.device)
        disps = torch.index_select(self.disps, 0, dirty_index)
        intrinsic = self.intrinsics[0].detach() * 1.0
        depths = 1.0 / disps
        thresh = self.mv_thresh * depths.mean(dim=[1, 2])
        count = droid_backends.depth_filter(self.poses, self.disps, intrinsic, dirty_index, thresh)
        multiview_masks = (count >= self.visible_num)
        depths[~multiview_masks] = torch.nan
        depths_median = depths.view(depths.shape[0], -1).nanmedian(dim=1).values
        masks = depths < 3 * depths_median[:, None, None]
        self.valid_depth_mask_small[dirty_index] = masks

    # :198-288
    def dspo(self, target, weight, eta, ii, jj, t0=1, t1=None, itrs=2, lm=1e-4, ep=0.1, motion_only=False,
             opt_type="pose_depth"):
        if t1 is None:
            t1 = max(ii.max().item(), jj.max().item()) + 1
        if opt_type == "pose_depth":
            target = target.view(-1, self.ht // self.down_scale, self.wd // self.down_scale, 2).permute(0, 3, 1, 2).contiguous()
            weight = weight.view(-1, self.ht // self.down_scale, self.wd // self.down_scale, 2).permute(0, 3, 1, 2).contiguous()
            droid_backends.ba(self.poses, self.disps, self.intrinsics[0], self.zeros, target, weight, eta, ii, jj, t0, t1,
                              itrs, lm, ep, motion_only, False)
            self.disps.clamp_(min=1e-5)
            self.calls.append("pose_depth")
            return True
        elif opt_type == "depth_scale":
            poses = SE3Stub(self.poses[None])
            disps = self.disps[None]
            scales, shifts = self.depth_scale, self.depth_shift
            ignore_frames = 0
            self.update_valid_depth_mask()
            curr_idx = self.counter_value - 1
            mono_d, est_d = self.mono_disps[:curr_idx + 1], self.disps[:curr_idx + 1]
            valid_d = self.valid_depth_mask_small[:curr_idx + 1]
            scale_t, shift_t, error_t = common.align_scale_and_shift(mono_d, est_d, valid_d)
            avg_disps = est_d.mean(dim=[1, 2])
            scales[:curr_idx + 1] = scale_t
            shifts[:curr_idx + 1] = shift_t
            target_t, weight_t, eta_t, ii_t, jj_t = target, weight, eta, ii, jj
            if self.mono_thres:
                invalid_mono_mask = (error_t / avg_disps > self.mono_thres) | (error_t.isnan()) | (scale_t < 0) | \
                                    (valid_d.sum(dim=[1, 2]) < valid_d.shape[1] * valid_d.shape[2] * 0.5)
                invalid_mono_index, = torch.where(invalid_mono_mask.clone())
                invalid_ii_mask = (ii < 0)
                idx_in_ii = torch.unique(ii)
                for idx in invalid_mono_index:
                    invalid_ii_mask = invalid_ii_mask | (ii == idx) | (jj == idx)
                target_t, weight_t = target[:, ~invalid_ii_mask], weight[:, ~invalid_ii_mask]
                ii_t, jj_t = ii[~invalid_ii_mask], jj[~invalid_ii_mask]
                idx_in_ii_t = torch.unique(ii_t)
                valid_eta_mask = torch.tensor([idx in idx_in_ii_t for idx in idx_in_ii]).to(self.device)
                eta_t = eta[valid_eta_mask]
            self.last_filter = (ii_t, jj_t, eta_t.shape[0])
            success = False
            for _ in range(itrs):
                if self.counter_value > ignore_frames and ii_t.shape[0] > 0:
                    poses, disps, wqs = geom_ba.BA_with_scale_shift(
                        target_t, weight_t, eta_t, poses, disps, self.intrinsics[None], ii_t, jj_t, self.mono_disps[None],
                        scales[None], shifts[None], self.valid_depth_mask_small[None], ignore_frames, lm, ep, alpha=0.01)
                    scales, shifts = wqs[0, :, 0], wqs[0, :, 1]
                    success = True
            self.depth_scale, self.depth_shift = scales, shifts
            self.disps = disps.squeeze(0)
            self.poses = poses.vec().squeeze(0)
            self.disps.clamp_(min=1e-5)
            self.calls.append("depth_scale" if success else "depth_scale(failed)")
            return success
        raise NotImplementedError

    # :290-298
    def ba(self, target, weight, eta, ii, jj, t0=1, t1=None, iters=2, lm=1e-4, ep=0.1, motion_only=False,
           opt_type="pose_depth"):
        if self.BA_type == "DSPO":
            success = self.dspo(target, weight, eta, ii, jj, t0, t1, iters, lm, ep, motion_only, opt_type)
            if not success:
                self.dspo(target, weight, eta, ii, jj, t0, t1, iters, lm, ep, motion_only, "pose_depth")
        elif self.BA_type == "DBA":
            self.dspo(target, weight, eta, ii, jj, t0, t1, iters, lm, ep, motion_only, "pose_depth")
        else:
            raise NotImplementedError

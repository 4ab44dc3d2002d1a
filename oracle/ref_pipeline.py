"""The reference's per-frame path assembled from its own CUDA extensions (oracle/_ref/*.so) and torch ops.

TEST / BASELINE INFRASTRUCTURE ONLY (bench.py --impl reference, end-to-end parity tests).  Nothing in
gaussianprediction_b200/ imports this.

The reference's Python glue cannot be imported (scene/gaussian_model.py needs tinycudann, pytorch3d, plyfile, ...
which are absent) so the train-iteration call sequence is restated here op for op, each step citing the line it
follows; every CUDA kernel that runs is the reference's own, unmodified, built for sm_100:
  render()                 gaussian_renderer/__init__.py:18-115
  GaussianModel.forward    scene/gaussian_model.py:231-304 (keypoint stage)
  get_nearest_mask         :110-125 -> frnn_grid_points (oracle/ref_cuda.py restates frnn.py:22-170)
  fill_nearest             :214-229     get_motion_delta :180-189     get_rotation_ :314-315
  _RasterizeGaussians      submodules/diff-gaussian-rasterization-w-depth/diff_gaussian_rasterization/__init__.py:40-132
The tinycudann blend-weight network is replaced on both arms by a logits tensor.
"""
import torch
import torch.nn as nn

from . import deform_oracle as do
from . import ref_cuda


class _RefRasterize(torch.autograd.Function):
    """diff_gaussian_rasterization/__init__.py:40-132 on the reference ``_C`` module."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, opacities, scales, rotations, cam, bg, sh_degree):
        fwd = ref_cuda.rasterize(means3D, opacities, bg, cam.world_view_transform, cam.full_proj_transform,
                                 cam.camera_center, cam.image_width, cam.image_height, cam.tanfovx, cam.tanfovy, shs=sh,
                                 scales=scales, rotations=rotations, sh_degree=sh_degree)
        ctx.fwd = fwd
        ctx.mark_non_differentiable(fwd["radii"], fwd["depth"], fwd["tidx"])
        return fwd["color"], fwd["radii"], fwd["depth"], fwd["tidx"]

    @staticmethod
    def backward(ctx, grad_out_color, _r, _d, _t):
        g = ref_cuda.rasterize_backward(ctx.fwd, grad_out_color.contiguous())
        return g["means3D"], g["means2D"], g["sh"], g["opacity"], g["scales"], g["rotations"], None, None, None


class RefKeypointModel(nn.Module):
    """The tensors and torch ops of the reference's keypoint-stage ``GaussianModel.forward`` (dense formulation)."""

    def __init__(self, args, xyz, rotation, scaling, opacity, motion_feature, super_gaussians, super_gaussians_feature,
                 logits, features_dc, features_rest, df_state=None):
        super().__init__()
        from gaussianprediction_b200.deform import DeformableField   # same layer stack as scene/deformable_field.py:74-127
        P = nn.Parameter
        self.args = args
        self._xyz, self._rotation, self._scaling, self._opacity = P(xyz), P(rotation), P(scaling), P(opacity)
        self.motion_feature = P(motion_feature)
        self.super_gaussians, self.super_gaussians_feature = P(super_gaussians), P(super_gaussians_feature)
        self.logits = P(logits)
        self._features_dc, self._features_rest = P(features_dc), P(features_rest)
        self.df_model = DeformableField(2 * args.time_freq + 6 * args.xyz_freq + args.feature_dim, 7, args.d, args.w)
        if df_state is not None:
            self.df_model.load_state_dict(df_state)
        self.active_sh_degree = 3
        # weights_fill: dense zeros [N, Kp_max] allocated once (scene/gaussian_model.py fill_nearest reads a slice)
        self.register_buffer("weights_fill", torch.zeros(xyz.shape[0], super_gaussians.shape[0]), persistent=False)

    @torch.no_grad()
    def get_nearest_mask(self):                                       # :110-125
        a = self.args
        if a.knn_type == "3D":
            _, nearest = ref_cuda.frnn_grid_points(self._xyz.unsqueeze(0), self.super_gaussians.unsqueeze(0), a.nearest_num, 1e3)
        else:
            concate_xyz = torch.cat([self._xyz, self.motion_feature * a.feature_amplify], dim=-1)
            concate_super_xyz = torch.cat([self.super_gaussians, self.super_gaussians_feature * a.feature_amplify], dim=-1)
            _, nearest = ref_cuda.frnn_grid_points(concate_xyz.unsqueeze(0), concate_super_xyz.unsqueeze(0), a.nearest_num, 1e8)
        self.nearest_mask = nearest.view([self._xyz.shape[0], a.nearest_num])

    def fill_nearest(self, weights):                                  # :214-229
        K = self.args.nearest_num
        weights_xyz = nn.Softmax(dim=-1)(weights[..., 0:K])
        weights_r = nn.Softmax(dim=-1)(weights[..., K:2 * K])
        fill = self.weights_fill[..., :self.super_gaussians.shape[0]]
        return torch.scatter(fill, -1, self.nearest_mask, weights_xyz), torch.scatter(fill, -1, self.nearest_mask, weights_r)

    def forward(self, t, iteration):                                  # :231-304, iteration > second_stage_iter
        a = self.args
        normalize = torch.nn.functional.normalize
        t = do.positional_encoding(t, a.time_freq).unsqueeze(0)                                   # :233
        factor = 0.1 * (1 - min(1, (iteration - a.second_stage_iteration) / a.xyz_noise_iteration))
        decay_noise = torch.randn_like(self.super_gaussians) * factor                              # :254
        xyz_embed = do.positional_encoding(self.super_gaussians + decay_noise, a.xyz_freq)         # :255
        weights = self.logits                                                                      # :257 (stand-in)
        self.get_nearest_mask()                                                                    # :260
        weights_xyz, weights_r = self.fill_nearest(weights)                                        # :261
        self.weights_sum = torch.abs(weights_xyz) + torch.abs(weights_r)                           # :263
        tt = t.repeat((self.super_gaussians_feature.shape[0], 1))                                  # :181
        delta = self.df_model(torch.cat([self.super_gaussians_feature, xyz_embed, tt], dim=-1))    # :183
        delta_xyz, delta_q = delta[..., 0:3], delta[..., 3:7]
        if a.norm_rotation:
            delta_q = normalize(delta_q)                                                           # :266-267
        delta_q = weights_r @ delta_q                                                              # :272
        delta_xyz = weights_xyz @ delta_xyz                                                        # :273
        xyz = self._xyz + delta_xyz                                                                # :285
        q = normalize(do.quat_mul(normalize(delta_q), self._rotation))                             # :286
        return xyz, q, torch.exp(self._scaling), torch.sigmoid(self._opacity)                      # :287, :304

    @torch.no_grad()
    def dense_weights(self, t, iteration):
        """``_, _, _, _, weights_xyz, weights_r = gaussians(time_, iteration, return_weights=True)`` (eval.py:125-126):
        the dense [N,Kp] matrices the extrapolation loop reuses for every frame."""
        self.get_nearest_mask()
        fill = self.weights_fill[..., :self.super_gaussians.shape[0]]
        K = self.args.nearest_num
        wx = torch.scatter(fill, -1, self.nearest_mask, nn.Softmax(dim=-1)(self.logits[..., 0:K]))
        wr = torch.scatter(fill, -1, self.nearest_mask, nn.Softmax(dim=-1)(self.logits[..., K:2 * K]))
        return wx, wr

    @property
    def get_features(self):                                           # :155-158
        return torch.cat((self._features_dc, self._features_rest), dim=1)


def ref_render(cam, pc, bg, time, it):
    """gaussian_renderer/__init__.py:18-115 with ``time`` given."""
    screenspace_points = torch.zeros_like(pc._xyz, requires_grad=True) + 0                         # :27
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    means3D, rotations, scales, opacity = pc(time, it)                                             # :74
    shs = pc.get_features                                                                          # :93
    color, radii, depth, tidx = _RefRasterize.apply(means3D, screenspace_points, shs, opacity, scales, rotations, cam,
                                                    bg, pc.active_sh_degree)                       # :98-106
    return {"render": color, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii,
            "depth": depth, "tidx": tidx}


@torch.no_grad()
def ref_render_motion(cam, pc, bg, weights_xyz, weights_r, kpts_i, kpts_rotation_i):
    """One frame of ``render_kpts`` (eval.py:137-153): the two dense matmuls + quaternion compose (:140-141), then
    ``render_motion`` (gaussian_renderer/__init__.py:117-191: canonical scales / opacity, the reference rasterizer)."""
    normalize = torch.nn.functional.normalize
    xyz_final = pc._xyz + (weights_xyz @ (kpts_i - pc.super_gaussians))                             # eval.py:140
    delta_r = normalize(do.quat_mul(normalize(weights_r @ kpts_rotation_i), pc._rotation))         # eval.py:141
    fwd = ref_cuda.rasterize(xyz_final, torch.sigmoid(pc._opacity), bg, cam.world_view_transform, cam.full_proj_transform,
                             cam.camera_center, cam.image_width, cam.image_height, cam.tanfovx, cam.tanfovy,
                             shs=pc.get_features, scales=torch.exp(pc._scaling), rotations=delta_r,
                             sh_degree=pc.active_sh_degree)
    return {"render": fwd["color"], "radii": fwd["radii"]}

"""Hot-path option names of the reference, with the values its shipped scripts use.

The reference configures everything through one mutable argparse namespace
``opt`` (options/base_options.py:268-287) that the hot-path classes re-read at
call time (``opt.is_train``, ``opt.query_size``, ``opt.prob``).  The drop-in
classes in this package take that same object; this module only provides a
namespace pre-filled with the per-scene optimisation values of
``dev_scripts/w_n360/lego.sh:40-101`` so that tests, ``bench.py`` and users
without the reference's option parser get an equivalent ``opt``.

Flag definitions: models/neural_points/neural_points.py:55-230,
models/aggregators/point_aggregators.py:15-217,
models/base_rendering_model.py:156-203,
models/neural_points_volumetric_model.py:47-52.
"""
from argparse import Namespace

# name -> value used by dev_scripts/w_n360/lego.sh (per-scene optimisation)
_LEGO = dict(
    # ---- querier / NeuralPoints (neural_points.py:55-230)
    K=8, SR=80, P=9, max_o=830000, NN=2,
    vsize=(0.004, 0.004, 0.004), vscale=(2, 2, 2),
    kernel_size=(3, 3, 3), query_size=(3, 3, 3),
    ranges=(-0.638, -1.141, -0.346, 0.634, 1.149, 1.141),
    z_depth_dim=400, radius_limit_scale=4.0, depth_limit_scale=0.0,
    wcoord_query=1, inverse=0, gpu_maxthr=1024, is_train=0, shpnt_jitter="uniform",
    point_features_dim=32, xyz_grad=0, feat_grad=1, conf_grad=1, dir_grad=1, color_grad=1,
    load_points=0, feature_init_method="rand", default_conf=0.15,
    # ---- aggregator (point_aggregators.py:15-217)
    which_agg_model="viewmlp", agg_distance_kernel="linear", agg_axis_weight=(1.0, 1.0, 1.0),
    agg_dist_pers=20, apply_pnt_mask=1, agg_intrp_order=2, agg_weight_norm=1,
    shading_feature_mlp_layer0=1, shading_feature_mlp_layer1=2, shading_feature_mlp_layer2=0,
    shading_feature_mlp_layer3=2, shading_alpha_mlp_layer=1, shading_color_mlp_layer=4,
    shading_feature_num=256, shading_color_channel_num=128, point_hyper_dim=256,
    dist_xyz_freq=5, num_feat_freqs=3, dist_xyz_deno=0.0, view_ori=0,
    agg_feat_xyz_mode="None", agg_alpha_xyz_mode="None", agg_color_xyz_mode="None",
    act_type="LeakyReLU", act_super=1, point_color_mode="1", point_dir_mode="1", point_conf_mode="1",
    weight_xyz_freq=2, weight_feat_dim=8, sh_degree=4, modulator_concat=0,
    # ---- rendering model (base_rendering_model.py:156-203, neural_points_volumetric_model.py:47-52)
    num_pos_freqs=10, num_viewdir_freqs=4, which_render_func="radiance", which_blend_func="alpha",
    which_tonemap_func="off", raydist_mode_unit=1,
    sparse_loss_weight=0.0, zero_one_loss_items=("conf_coefficient",), prob=0,
)


def make_opt(**overrides) -> Namespace:
    """Namespace with the reference's hot-path attribute names (lego.sh values)."""
    d = dict(_LEGO)
    unknown = set(overrides) - set(d)
    if unknown:
        raise KeyError("not a hot-path option of the reference: %s" % sorted(unknown))
    d.update(overrides)
    return Namespace(**d)

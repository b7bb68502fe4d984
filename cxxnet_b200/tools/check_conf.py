"""Report configuration keys that nothing reads.

    python -m cxxnet_b200.tools.check_conf net.conf [name=value ...]

A configuration is a flat list of pairs and every consumer silently ignores what it does not know (reference
``src/cxxnet_main.cpp:200-202``), so a misspelt key - the reference's own examples carry ``train_eval`` for
``eval_train`` and ``slient`` for ``silent`` - changes nothing and says nothing.  This lists every such key, the layer or iterator it was written under, and
the nearest known key.  Exit status 1 when something was found.
"""
from __future__ import annotations

import difflib
import re
import sys

from ..config import parse_cli_overrides, parse_config_file

TASK = """task dev model_in model_dir continue start_counter num_round max_round save_model print_step silent test_io
extract_node_name output_format net_type reset_net_type data eval pred iter""".split()
TRAINER = """batch_size update_period eval_train seed param_server metric backend cuda_graph test_on_server layer_timing step_trace
input_shape updater sync netconfig extra_data_num label_vec extra_data_shape check_nan
pull_at_backprop bigarray_bound update_on_server num_servers init_on_worker force_contiguous fullc_gather fullc_shard""".split()
LAYER = """init_sigma init_uniform init_bias init_sparse random_type nhidden nchannel ngroup kernel_size kernel_height
kernel_width stride pad pad_y pad_x no_bias temp_col_max algo threshold local_size alpha beta knorm b lb ub calm_start
calm_end keep init_slope random_slope random eps fixconn_weight target grad_scale""".split()
UPDATER = """lr eta wd momentum momentum_schedule base_momentum final_momentum saturation_epoch clip_gradient beta1 beta2
lr:schedule lr:gamma lr:alpha lr:step lr:factor lr:minimum_lr lr:start_epoch
eta:schedule eta:gamma eta:alpha eta:step eta:factor eta:minimum_lr eta:start_epoch""".split()
ITER = """path_img path_label input_flat shuffle index_offset seed_data synthetic num_synthetic image_list image_root
image_bin image_conf_prefix image_conf_ids dist_num_worker dist_worker_rank label_width native_reader
source_shape num_inst num_class value_scale
rand_crop rand_mirror mirror crop_x_start crop_y_start divideby scale image_mean mean_value max_random_contrast
max_random_illumination max_rotate_angle max_shear_ratio max_aspect_ratio min_crop_size max_crop_size min_random_scale
max_random_scale min_img_size max_img_size fill_value rotate rotate_list warp_interp
round_batch test_skipread input_dtype native_pipeline decode_threads slice_world slice_rank buffer_size max_nbatch
filename""".split()
KNOWN = set(TASK + TRAINER + LAYER + UPDATER + ITER)


def known(name: str) -> bool:
    if name in KNOWN:
        return True
    if re.match(r"(layer|metric|label_vec|extra_data_shape)\[", name):
        return True
    for prefix in ("wmat:", "bias:", "master:", "slave:"):
        if name.startswith(prefix):
            return known(name[len(prefix):])
    return False


def check(pairs):
    """[(line_or_None, name, context, suggestion)] for the pairs nothing reads"""
    out = []
    context = ""
    for item in pairs:
        name, val = item[0], item[1]
        if name in ("data", "eval", "pred"):
            context = "%s block" % name
        elif name == "iter":
            context = "" if val == "end" else "iter = %s" % val
        elif name.startswith("layer["):
            context = "%s = %s" % (name, val)
        elif name == "netconfig" and val == "end":
            context = ""
        if not known(name):
            bare = name.split(":")[-1]
            near = difflib.get_close_matches(bare, sorted(KNOWN), n=1, cutoff=0.75)
            if not near:                     # the same words in another order (train_eval / eval_train)
                near = [k for k in sorted(KNOWN) if sorted(k.split("_")) == sorted(bare.split("_"))][:1]
            out.append((name, val, context, near[0] if near else ""))
    return out


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    if not argv:
        print(__doc__)
        return 0
    pairs = list(parse_config_file(argv[0])) + list(parse_cli_overrides(argv[1:]))
    found = check(pairs)
    for name, val, context, near in found:
        print("%s = %s   <- no consumer%s%s" % (name, val, (" (under %s)" % context) if context else "",
                                                (";  did you mean `%s`?" % near) if near else ""))
    print("%d pairs, %d without a consumer" % (len(pairs), len(found)))
    return 1 if found else 0


if __name__ == "__main__":
    sys.exit(main())

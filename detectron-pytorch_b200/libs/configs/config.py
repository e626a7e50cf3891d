"""The hot-path keys of the reference's module-as-namespace config (libs/configs/config.py:11-172).
Like upstream they are plain module globals that the layer functions read AT CALL TIME
(libs/layers/anchor.py:50-52,60,86,95,111), so `cfg.rpn_fg_threshold = 0.7` behaves as it does there.
Defaults are the reference's."""
num_classes = 81                                  # config.py:65
class_activation = 'sigmoid'                      # config.py:67
canvas_width = 1024
canvas_height = 1024
strides = (8, 16, 32, 64, 128)                    # config.py:82
rpn_bg_threshold = 0.4                            # config.py:98
rpn_fg_threshold = 0.6
rpn_batch_size = 384
rpn_fg_fraction = 0.25
rpn_clobber_positives = True
rpn_sample_strategy = 'simple'                    # 'simple' | 'traditional' (np.random draws in the reference's order)
rpn_box_encoding = 'fastrcnn'                     # 'linear', 'fastrcnn'
anchor_scales = [[2, 2.52, 3.17], [4, 5.04, 6.35], [8, 10.08, 12.70], [16, 20.16, 25.40], [32, 40.32, 50.80]]
anchor_ratios = [0.5, 1, 2.0]
anchor_base = 16
anchor_shift = [[0.0, 0.0], ]
bbweights = [1., 1., 1., 1.]                      # config.py:33
fg_threshold = 0.5
rois_per_image = 128
roi_fg_fraction = 0.25
roi_bg_threshold = 0.3
max_det_num = 100
score_threshold = 0.5
overlap_threshold = 0.5
overlap_threshold_noncrowd = 0.4                  # config.py:151
noncrowd_classes = [5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 16] + list(range(25, 34)) + list(range(36, 56)) + list(range(57, 81))  # config.py:154-158
batch_size = 6                                    # config.py:78
ignored_area_intersection_fraction = 0.3


def load_from_yaml(path):
    """libs/configs/config.py:176-202 restricted to the keys above (type-checked overlay)."""
    import yaml
    g = globals()
    with open(path) as f:
        for k, v in (yaml.safe_load(f) or {}).items():
            if k in g and not callable(g[k]):
                if g[k] is not None and not isinstance(v, type(g[k])) and not (isinstance(g[k], (tuple, list)) and isinstance(v, (tuple, list))) \
                        and not (isinstance(g[k], float) and isinstance(v, int)):
                    raise ValueError("type mismatch for config key %s: %r vs %r" % (k, type(v), type(g[k])))
                g[k] = v

"""Host logic of wheat_3dgs_b200.train.Trainer that needs no GPU: the position learning-rate schedule and the SH-degree
steps of the reference's loop head (train_vanilla_3dgs.py:58-62).

The expected rates are the reference's own: utils/general_utils.py get_expon_lr_func called as
scene/gaussian_model.py:182-186 calls it (lr_init = 0.00016 * s, lr_final = 0.0000016 * s, lr_delay_mult = 0.01,
max_steps = 30000; lr_delay_steps keeps its default 0), evaluated in this container with
    f = get_expon_lr_func(lr_init=0.00016 * s, lr_final=0.0000016 * s, lr_delay_mult=0.01, max_steps=30000); f(it)
for s = 1, and with lr_delay_steps=500 for s = 4.2 (the ease-in branch the default never takes)."""
import torch

from wheat_3dgs_b200 import train

REF_DEFAULT = {0: 0.00016, 1: 0.0001599754409773643, 1000: 0.00013723134377454326, 7000: 5.463278198133772e-05,
               30000: 1.5999999999999995e-06, 40000: 1.5999999999999995e-06}
REF_DELAY_500_SCALE_4_2 = {1: 8.808683036000275e-06, 250: 0.0004591798266929255, 1000: 0.0005763716438530808,
                           30000: 6.72e-06}


def _trainer(deg=0, **kw):
    P = 8
    raw = {"xyz": torch.zeros(P, 3), "f_dc": torch.zeros(P, 1, 3), "f_rest": torch.zeros(P, 15, 3),
           "opacity": torch.zeros(P, 1), "scaling": torch.zeros(P, 3), "rotation": torch.ones(P, 4)}
    return train.Trainer(raw, deg, "cpu", **kw)


def test_position_lr_follows_the_reference_schedule():
    t = _trainer()
    assert t.get_lr("xyz") == 0.00016
    for it, want in REF_DEFAULT.items():
        assert t.update_learning_rate(it) == want          # same float64 operations: equal, not close
        assert t.get_lr("xyz") == want
        assert t.get_lr("opacity") == 0.05 and t.get_lr("f_rest") == 0.0025 / 20   # only xyz is scheduled
    assert train.expon_lr(-1, 1e-3, 1e-5) == 0.0 and train.expon_lr(5, 0.0, 0.0) == 0.0


def test_position_lr_with_extent_scale_and_ease_in():
    t = _trainer(spatial_lr_scale=4.2, position_lr=dict(train.POSITION_LR, lr_delay_steps=500))
    for it, want in REF_DELAY_500_SCALE_4_2.items():
        assert t.update_learning_rate(it) == want


def test_set_lr_reaches_the_optimiser_group():
    t = _trainer()
    t.set_lr("scaling", 0.25)
    assert [g["lr"] for g in t.opt.param_groups if g["name"] == "scaling"] == [0.25]


def test_sh_degree_rises_every_1000_iterations_up_to_the_coefficient_count():
    t = _trainer(deg=0)
    assert t.max_sh_degree == 3 and t.active_sh_degree == 0
    for _ in range(5):
        t.oneup_sh_degree()
    assert t.active_sh_degree == 3

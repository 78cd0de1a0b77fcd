"""GPU parity of the lambdarank objective (all 18 lambdarank_target values of the fork) end to end against the
reference CPU objective: identical trees and scores after several boosting iterations."""
import numpy as np
import pytest

from lambdagap_b200 import engine
from oracle import ref_lgbm as R
from tests import helpers as H
from tests.test_booster_parity import _compare_models, _train_both

pytestmark = pytest.mark.gpu

TARGETS = ["ndcg", "lambdaloss-ndcg", "lambdaloss-ndcg-plus-plus", "bndcg", "lambdaloss-bndcg",
           "lambdaloss-bndcg-plus-plus", "precision", "arpk", "lambdaloss-arp1", "lambdaloss-arp2", "ranknet",
           "bin-ranknet", "lambdagap-s", "lambdagap-x", "lambdagap-s-plus", "lambdagap-x-plus", "lambdagap-s-plus-plus",
           "lambdagap-x-plus-plus"]


@pytest.mark.parametrize("target", TARGETS)
def test_lambdarank_target(gpu, target):
    # truncation level 5 <= every query length (the reference's `precision` nth_element is undefined beyond cnt)
    X, y, group = H.make_ranking(400, 12, seed=31, min_len=6, max_len=80)
    params = ("objective=lambdarank lambdarank_target=%s lambdarank_truncation_level=5 lambdagap_weight=1.5 "
              "num_leaves=15 min_data_in_leaf=10 learning_rate=0.1" % target)
    rb, ob = _train_both(X, y, params, 6, group=group)
    # fp32 lambda accumulation is replayed in the CPU's order; sum_lambdas is a block reduction -> tolerance case
    _compare_models(rb, ob, rtol=1e-5)
    np.testing.assert_allclose(ob.train_scores(raw=True), rb.train_scores(), rtol=1e-5, atol=1e-7)


def test_lambdarank_long_queries_and_weights(gpu):
    X, y, group = H.make_ranking(30, 8, seed=7, min_len=150, max_len=1500)
    w = np.random.default_rng(3).uniform(0.5, 1.5, size=len(y)).astype(np.float32)
    params = "objective=lambdarank lambdarank_truncation_level=30 num_leaves=31 min_data_in_leaf=20"
    rb, ob = _train_both(X, y, params, 4, group=group, weight=w)
    _compare_models(rb, ob, rtol=1e-5)

"""CPU: truth-recovery runs of scaled-down BASELINE configurations on the oracle and the golden density vectors through the
oracle's table interpreter (tests/outside_checks.py).  The GPU runs the same checks at larger sizes (test_gpu_outside_checks.py)."""
import time

import pytest

from tests import outside_checks as oc


def test_golden_vectors_through_the_table_path(oracle):
    # the same 592 vectors test_oracle_densities.py feeds to the scalar functions, here through model lowering,
    # parameter programs and calculate_nodes
    oc.check_golden_vectors(oracle, 5e-14)


def test_logistic_recovers_mle(oracle):
    oc.check_logistic(oracle, N=3000, p=4, chains=4, niter=5000)


def test_lmm_recovers_truth(oracle):
    oc.check_lmm(oracle, N=365, G=120, chains=4, niter=3000)   # last group larger: SEL_SEG, not SEL_STRIDE


def test_lmm_permuted_groups_recover_truth(oracle):
    oc.check_lmm(oracle, N=360, G=90, chains=4, niter=3000, permute=True)


def test_occupancy_recovers_truth(oracle):
    oc.check_occupancy(oracle, S=1500, J=4, chains=4, niter=2500)


def test_glmm_recovers_truth(oracle):
    oc.check_glmm(oracle, N=600, G=150, chains=4, niter=4000)   # equal groups: SEL_STRIDE


def test_glmm_unequal_groups_recover_truth(oracle):
    oc.check_glmm(oracle, N=607, G=150, chains=4, niter=4000)   # SEL_SEG


@pytest.mark.parametrize("which", ["seg", "stride"])
def test_a_broken_selection_is_caught(oracle, monkeypatch, which):
    """The reason these checks exist: shift a dependency selection by one node (both back ends would consume the same wrong
    table, so GPU-vs-oracle parity stays green) and the closed-form conditional means must object."""
    import numpy as np
    from nimble_b200 import table as T
    from nimble_b200.mcmc import _Lowering
    good = _Lowering._selection

    def broken(self, tgt, n, Tn):
        sel = good(self, tgt, n, Tn)
        if which == "seg" and sel[0] == T.SEL_SEG:
            off = np.minimum(self.b.iarr[sel[1]] + 1, n)
            off[0] = 0
            return (T.SEL_SEG, self.b.add_iarr(off), 0)
        if which == "stride" and sel[0] == T.SEL_STRIDE and sel[1] > 1:
            return (T.SEL_STRIDE, sel[1] - 1, 0)
        return sel
    monkeypatch.setattr(_Lowering, "_selection", broken)
    with pytest.raises(AssertionError):
        if which == "seg":
            oc.check_lmm(oracle, N=365, G=120, chains=4, niter=3000)
        else:
            oc.check_glmm(oracle, N=600, G=150, chains=4, niter=4000)

"""The bf16 native engine against the unmodified fp32 reference over several SalientGrads rounds on the same GPU, same data, same
initial weights, dropout off, 15 % label noise (scripts/convergence_vs_reference.py): the saliency mask agrees up to near-ties (a few per cent of the entries), the
global test loss follows the reference's round by round, the final accuracy is within two test samples."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_native_bf16_engine_follows_the_fp32_reference_trajectory():
    sys.path.insert(0, ROOT)
    from baseline import ref_harness
    why = ref_harness.reference_available()
    if why:
        pytest.skip(why)
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import convergence_vs_reference as cv
    res = cv.run(rounds=6)
    # measured 0.056 at 6-10 rounds: bf16 activations move near-tie saliency scores across the top-k threshold
    assert res["mask_hamming_fraction"] < 0.08, res["mask_hamming_fraction"]
    assert res["max_abs_loss_diff"] < 0.05, (res["reference"]["loss"], res["ours"]["loss"])
    assert res["final_acc_diff"] <= 2 * res["one_test_sample_in_acc"] + 1e-9, (res["reference"]["acc"], res["ours"]["acc"])
    # the task is not trivially separable: nobody reaches 100 %
    assert max(res["reference"]["acc"]) < 1.0 and max(res["ours"]["acc"]) < 1.0

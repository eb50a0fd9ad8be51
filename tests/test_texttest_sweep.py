"""The reference's own texttest scenarios through the engine, in bulk (tools_dev/texttest_sweep.py; results of the full sweep over 4585
test directories in profiles/r02_texttest_sweep.tsv: every scenario the engine accepts -- 561 -- reproduces the reference's trajectories
bit for bit, everything else is refused with a message, rejected by the reference itself, or has no vehicles).

This test re-runs every eighth of the accepted scenarios: the reference's loader flattens the scenario and records the reference's
own run (oracle/_ref/sumo_flatten), tests/host_emul (the device functions compiled for the host) must give byte-identical traces.
It also checks that a handful of scenarios with features outside the engine's scope are still REFUSED (never simplified).
Needs /root/reference (the build container); skipped where it is absent."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools_dev"))
import texttest_sweep  # noqa: E402

SWEEP = os.path.join(ROOT, "profiles", "r02_texttest_sweep.tsv")
SAME = [l.split("\t")[1] for l in open(SWEEP) if l.startswith("SAME\t")]
REFUSED = {
    "bugs/0xxx/ticket720": "vaporizers are not supported",
    "extended/bidiEdge/1edge_2vehicles": "bidirectional lanes",
    "extended/state/load_remove": "--load-state is not supported",
    "cf_model/Krauss/sigma1_step0.1_sigmaStep1": "sigmaStep longer than the step length",
    "spec/departure/insertionChecks/none": "insertionChecks other than 'all'",
    "spec/departure/errors/slow_lane_ahead_ignore": "--ignore-route-errors is not supported",
    "opposite_direction_driving/bugs/ticket9721": "opposite-direction driving",
    "rail/slow_fast_slow": "only static programs are supported",
    "output/meandata_edgebased/jam1": "variableSpeedSigns are not supported",
    "bugs/12xxx/ticket12721c": "option --collision.action is not modelled",
    "junction_model/opposite_left_turn/small_junction": "option --collision.check-junctions is not modelled",
}


def need_reference(host_emul):
    if not os.path.isdir(texttest_sweep.TT_ROOT):
        pytest.skip("reference sources (tests/sumo) are not available on this machine")
    texttest_sweep.EMUL = host_emul
    os.makedirs("/tmp/w", exist_ok=True)


@pytest.mark.parametrize("rel", SAME[::8])
def test_reference_texttest_scenario_is_reproduced_bit_for_bit(rel, host_emul):
    need_reference(host_emul)
    _, status, info = texttest_sweep.run_one(rel)
    assert status == "SAME", (rel, status, info)


@pytest.mark.parametrize("rel", sorted(REFUSED))
def test_reference_texttest_scenario_outside_the_scope_is_refused(rel, host_emul):
    need_reference(host_emul)
    _, status, info = texttest_sweep.run_one(rel)
    assert status == "REFUSED" and REFUSED[rel] in info, (rel, status, info)


def test_sweep_file_has_no_divergence():
    text = open(SWEEP).read()
    assert "\nDIFF\t" not in text and "\nFLOWDIFF\t" not in text and len(SAME) >= 560

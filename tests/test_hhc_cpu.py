"""CPU tests of the heavy-hexagon emitter (qiskit_qec_b200/hhc.py): sizes and
labels follow reference circuits/hhc_circuit.py:12-53,98-103,474-489; and the emitted schedule
MEASURES what the reference's does -- through the C oracle's propagate, an X (Z) error on a data
qubit flips exactly the Z (X) gauge outcomes that contain the qubit in every later round, no
flag, and the qubit's own readout bit in the matching basis."""

import numpy as np
import pytest

from qiskit_qec_b200.hhc import HHCCode, hex_ancillas, hhc_circuit

OPTIONS = [(3, {}), (3, dict(group_meas=True)), (5, dict(blocks=2)),
           (3, dict(round_schedule="xz", basis="x", blocks=2)),
           (3, dict(num_initialize=1, idle_before_measure=True)),
           (3, dict(num_initialize=0, init_error=True, barriers=False, logical_paulis="xz"))]


def test_code_geometry_and_ancillas():
    code = HHCCode(3)
    assert code.x_gauges == [[0, 3], [1, 4], [2, 5], [3, 6], [4, 7], [5, 8]]
    assert code.z_gauges == [[0, 1], [1, 2, 4, 5], [3, 4, 6, 7], [7, 8]]
    total, xa, za = hex_ancillas(code)
    assert total == 14 and xa == [[9], [10], [11], [12], [13], [14]]
    # boundary gauges own three ancillas; bulk ones borrow the X-gauge ancillas left and right
    assert za == [[15, 16, 17], [10, 18, 11], [12, 19, 13], [20, 21, 22]]
    with pytest.raises(ValueError):
        HHCCode(4)


@pytest.mark.parametrize("d,kw", OPTIONS)
def test_sizes_and_labels(d, kw):
    c = hhc_circuit(d=d, **kw)
    code = HHCCode(d)
    xg, zg = len(code.x_gauges), len(code.z_gauges)
    sched, blocks = kw.get("round_schedule", "zx"), kw.get("blocks", 1)
    assert c.nc == blocks * sum(xg if r == "x" else 3 * zg for r in sched) + code.n
    assert c.nq == code.n + hex_ancillas(code)[0] and c.creg_sizes == [c.nc]
    counts = c.gate_counts()
    assert counts["measure"] == c.nc
    # every measure writes its own clbit, in increasing order within a round's groups
    assert sorted(row[2] for row in c.ops if row[0] == 8) == list(range(c.nc))
    labels = set(c.labels)
    assert ("idm" in labels) == kw.get("distinct_measurement_idle", True)
    assert ("idinit" in labels) == (kw.get("num_initialize") == 0)
    idbm = {l for l in labels if l.startswith("idbm")}
    assert idbm == ({f"idbm{k}" for k in range(c.nc)} if kw.get("idle_before_measure") else set())
    # a labelled idle lowers to the id opcode; its label is its noise channel
    assert all(c.ops[i][0] == 6 for i, l in enumerate(c.labels) if l.startswith("id"))


def test_option_validation_messages():
    for kw, msg in [(dict(blocks=0), "positive integer rounds"), (dict(basis="y"), "basis to be"),
                    (dict(initial_state="0"), "initial state"), (dict(round_schedule="zxy"), "round schedule"),
                    (dict(logical_paulis="x"), "len\\(logical paulis\\)"),
                    (dict(xprs=True, num_initialize=1), 'op "xprs" not recognized')]:
        with pytest.raises(Exception, match=msg):
            hhc_circuit(**kw)


@pytest.mark.parametrize("d,kw", OPTIONS[:5])
def test_schedule_measures_the_gauge_operators(port, d, kw):
    c, code = hhc_circuit(d=d, **kw), HHCCode(d)
    xg, zg = len(code.x_gauges), len(code.z_gauges)
    sched, blocks, basis = kw.get("round_schedule", "zx"), kw.get("blocks", 1), kw.get("basis", "z")
    prep = {}   # last preparation op (reset / h) of each data qubit before the first round
    for i, row in enumerate(c.ops[:2 * c.nq + 5]):
        if row[0] in (7, 0) and row[1] < code.n:
            prep[row[1]] = i
    for q in range(code.n):
        for pauli in "xz":
            got = port.propagate_batch(c.nq, c.nc, c.ops, [([prep[q]], [pauli])])[0]
            want, start = np.zeros(c.nc, np.uint8), 0
            for _ in range(blocks):
                for rs in sched:
                    gauges = code.x_gauges if rs == "x" else code.z_gauges
                    if (rs == "x") == (pauli == "z"):
                        for j, g in enumerate(gauges):
                            want[start + j] = q in g
                    start += xg if rs == "x" else 3 * zg
            want[start + q] = (basis == "z") == (pauli == "x")
            assert np.array_equal(got, want), (q, pauli)

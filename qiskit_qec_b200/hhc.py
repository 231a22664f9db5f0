"""Heavy-hexagon subsystem-code syndrome-extraction circuits as gate streams.

qiskit-free emitter of what the reference's ``HHCCircuit(...).syndrome_measurement()`` appends
(reference ``circuits/hhc_circuit.py:56-103`` options and labels, ``:151-241`` X-gauge round,
``:243-424`` Z-gauge round with flags, ``:426-574`` the driver), on the code geometry of
``codes/hhc.py:16-183``.  Output: an ``EncodedCircuit`` in the reference's append order, with the
LABELLED idles this fault-path machinery was written for -- ``idm`` (data idle during
measurement), ``idbm<clbit>`` (idle before a measurement), ``idinit`` (initialisation error) --
each its own noise channel through ``node_name_label`` (``utils/dag.py:18-27``).

The ``xprs`` option emits a custom gate the propagator rejects (``cerrorpropagator.py:67-68``:
``op "xprs" not recognized``); it is refused here with the same message.
"""

from __future__ import annotations

from typing import List, Optional, Tuple

from qiskit_qec_b200.circuits import EncodedCircuit

IDLE_MEASURE = "idm"
IDLE_BEFORE_MEASURE = "idbm"
IDLE_INIT = "idinit"


class HHCCode:
    """Heavy-hexagon compass code data (reference ``codes/hhc.py:16-183``): qubits numbered row by
    row; X gauges = vertical pairs, Z gauges = weight-4 faces (weight 2 on the top / bottom rows)."""

    def __init__(self, distance: int):
        if distance % 2 == 0:
            raise ValueError("even distance not implemented")
        if distance < 3:
            raise ValueError("require distance > 2")
        d = self.d = distance
        self.n = d * d
        at = lambda r, c: r * d + c  # noqa: E731
        self.x_gauges = [[at(i, j), at(i + 1, j)] for i in range(d - 1) for j in range(d)]
        self.z_gauges: List[List[int]] = []
        for i in range(d + 1):
            for j in range(d - 1):
                if i % 2 != j % 2:
                    continue
                if i == 0:
                    self.z_gauges.append([at(0, j), at(0, j + 1)])
                elif i == d:
                    self.z_gauges.append([at(d - 1, j), at(d - 1, j + 1)])
                else:
                    self.z_gauges.append([at(i - 1, j), at(i - 1, j + 1), at(i, j), at(i, j + 1)])
        self.logical_x = [list(range(self.n))]
        self.logical_z = [list(range(self.n))]


def hex_ancillas(code: HHCCode) -> Tuple[int, List[List[int]], List[List[int]]]:
    """(total ancillas, per X gauge [ancilla], per Z gauge [left, middle, right]) as absolute
    qubit indices (reference ``hhc_circuit.py:12-53``): one ancilla per X gauge; a boundary Z
    gauge gets three of its own, a bulk Z gauge borrows the ancillas of the X gauges on its left
    and right columns and adds one in the centre of the face."""
    x_anc = [[code.n + j] for j in range(len(code.x_gauges))]
    total = len(code.x_gauges)
    z_anc = []
    for zg in code.z_gauges:
        if len(zg) == 2:
            z_anc.append([code.n + total, code.n + total + 1, code.n + total + 2])
            total += 3
            continue
        left = right = None
        for k, xg in enumerate(code.x_gauges):
            if zg[0] in xg and zg[2] in xg:
                left = code.n + k
            if zg[1] in xg and zg[3] in xg:
                right = code.n + k
        z_anc.append([left, code.n + total, right])
        total += 1
    return total, x_anc, z_anc


class _Emitter:
    def __init__(self, d, barriers, idles, distinct_measurement_idle, init_error, group_meas, xprs,
                 blocks, round_schedule, basis, initial_state, logical_paulis, num_initialize,
                 idle_before_measure):
        # the reference's _validate (:104-149), same conditions
        if blocks < 1:
            raise Exception("expected positive integer rounds")
        if set(round_schedule) > set("xz"):
            raise Exception("expected round schedule of 'x', 'z' chars")
        if basis not in ("x", "z"):
            raise Exception("expected basis to be 'x' or 'z'")
        if initial_state not in ("+", "-"):
            raise Exception("expected initial state '+' or '-'")
        if set(logical_paulis) > set("ixyz"):
            raise Exception("expected 'i', 'x', 'y', 'z' logical paulis")
        if len(logical_paulis) > 0 and len(logical_paulis) != len(round_schedule):
            raise Exception("len(logical paulis) != len(round schedule)")
        if num_initialize is not None and num_initialize < 0:
            raise Exception("expected zero or positive integer number of initialization")
        if xprs and num_initialize is not None:
            raise Exception('op "xprs" not recognized')
        self.code = HHCCode(d)
        self.barriers, self.idles, self.dmi = barriers, idles, distinct_measurement_idle
        self.init_error, self.group_meas, self.blocks = init_error, group_meas, blocks
        self.round_schedule, self.basis, self.initial_state = round_schedule, basis, initial_state
        self.logical_paulis, self.num_init, self.ibm = logical_paulis, num_initialize, idle_before_measure
        self.total_ancilla, self.x_anc, self.z_anc = hex_ancillas(self.code)
        self.nq = self.code.n + self.total_ancilla
        self.circ: Optional[EncodedCircuit] = None

    # ---- small pieces shared by the two rounds
    def _measure(self, q: int, c: int, final: bool):
        if self.ibm:
            self.circ.id(q, label=IDLE_BEFORE_MEASURE + str(c))
        self.circ.measure(q, c)

    def _reset_after(self, qubits, final: bool):
        if not final and self.num_init is not None:
            for q in qubits:
                self.circ.reset(q)

    def _logical_pauli(self, pauli):
        if pauli in ("x", "y"):
            for k in self.code.logical_x[0]:
                self.circ.x(k)
        if pauli in ("z", "y"):
            for k in self.code.logical_z[0]:
                self.circ.z(k)

    def _grouped_prologue(self, pauli, final: bool):
        """what precedes a grouped measurement (:207-217, :378-388)"""
        self._logical_pauli(pauli)
        if self.barriers and pauli is not None:
            self.circ.barrier()
        if self.basis == "x" and final:
            for i in range(self.code.n):
                self.circ.h(i)
        if not self.barriers:
            self.circ.barrier()

    def _data_idle_during_measurement(self, qubits, final: bool):
        for i in qubits:
            if self.dmi:
                if not self.group_meas or not final:
                    self.circ.id(i, label=IDLE_MEASURE)
            else:
                self.circ.id(i)

    # ---- X gauges: one ancilla each (:151-241)
    def x_round(self, cbits: List[int], final: bool, pauli):
        c, code = self.circ, self.code
        for anc in self.x_anc:
            for a in anc:
                if self.num_init is None:
                    c.reset(a)
                c.h(a)
        if self.barriers and self.num_init is None:
            c.barrier()
        for step in range(2):
            idle = list(range(code.n))
            for i, xg in enumerate(code.x_gauges):
                c.cx(self.x_anc[i][0], xg[step])
                idle.remove(xg[step])
            if self.idles:
                for i in idle:
                    c.id(i)
            if self.barriers:
                c.barrier()
        for i, anc in enumerate(self.x_anc):
            for a in anc:
                c.h(a)
                if not self.group_meas:
                    self._measure(a, cbits[i], final)
                    self._reset_after([a], final)
        if self.group_meas:
            self._grouped_prologue(pauli, final)
            for i, anc in enumerate(self.x_anc):
                for a in anc:
                    self._measure(a, cbits[i], final)
                    self._reset_after([a], final)
        if self.idles:
            self._data_idle_during_measurement(range(code.n), final)
        if self.barriers or (self.group_meas and not final):
            c.barrier()

    # ---- Z gauges with left / right flags: seven steps (:243-424; schedule of Fig. 2b of
    # arXiv:1907.09528 conjugated by Hadamards)
    def z_round(self, cbits: List[int], lflags: List[int], rflags: List[int], final: bool, pauli):
        c, code = self.circ, self.code
        none_init = self.num_init is None
        for step in range(7):
            idle = list(range(code.n))
            for j, zg in enumerate(code.z_gauges):
                left, middle, right = self.z_anc[j]
                boundary, top = len(zg) == 2, zg[0] < code.d
                lower_or_bulk = (boundary and not top) or not boundary
                if step == 0:
                    if none_init:
                        c.reset(middle)
                        c.reset(right)
                    c.h(right)
                elif step == 1:
                    c.cx(right, middle)
                    if none_init:
                        c.reset(left)
                    c.h(left)
                elif step == 2:
                    c.cx(left, middle)
                    if lower_or_bulk:
                        c.cx(zg[1], right)
                        idle.remove(zg[1])
                    elif self.idles:
                        c.id(right)
                elif step == 3:
                    if boundary and top:
                        pair = (zg[0], zg[1])
                    elif not boundary:
                        pair = (zg[2], zg[3])
                    else:
                        pair = None
                    if pair:
                        c.cx(pair[0], left)
                        c.cx(pair[1], right)
                        idle.remove(pair[0])
                        idle.remove(pair[1])
                    elif self.idles:
                        c.id(left)
                        c.id(right)
                elif step == 4:
                    c.cx(right, middle)
                    if lower_or_bulk:
                        c.cx(zg[0], left)
                        idle.remove(zg[0])
                    elif self.idles:
                        c.id(left)
                elif step == 5:
                    c.cx(left, middle)
                    if self.group_meas:
                        if self.idles:
                            c.id(right)
                    else:
                        c.h(right)
                        self._measure(right, rflags[j], final)
                        self._reset_after([right], final)
                else:
                    c.h(left)
                    if self.group_meas:
                        c.h(right)
                    else:
                        if self.ibm:
                            c.id(middle, label=IDLE_BEFORE_MEASURE + str(cbits[j]))
                            c.id(left, label=IDLE_BEFORE_MEASURE + str(lflags[j]))
                        c.measure(middle, cbits[j])
                        c.measure(left, lflags[j])
                        self._reset_after([middle, left], final)
            if step == 6 and self.group_meas:
                self._grouped_prologue(pauli, final)
                for j in range(len(code.z_gauges)):
                    left, middle, right = self.z_anc[j]
                    if self.ibm:
                        c.id(right, label=IDLE_BEFORE_MEASURE + str(rflags[j]))
                        c.id(middle, label=IDLE_BEFORE_MEASURE + str(cbits[j]))
                        c.id(left, label=IDLE_BEFORE_MEASURE + str(lflags[j]))
                    c.measure(right, rflags[j])
                    c.measure(middle, cbits[j])
                    c.measure(left, lflags[j])
                    self._reset_after([middle, left, right], final)
            if self.idles:
                if step == 6:
                    self._data_idle_during_measurement(idle, final)
                elif step != 0 or none_init:
                    for i in idle:
                        c.id(i)
            if self.barriers or (step == 6 and self.group_meas and not final):
                if step != 0 or none_init:
                    c.barrier()

    # ---- the driver (:426-574)
    def build(self) -> EncodedCircuit:
        code = self.code
        xg, zg = len(code.x_gauges), len(code.z_gauges)
        per_block = sum(xg if r == "x" else 3 * zg for r in self.round_schedule)
        nc = per_block * self.blocks + code.n
        c = self.circ = EncodedCircuit(nq=self.nq, nc=nc, name="hhc")
        c.creg_sizes = [nc]
        if self.num_init is None:
            for i in range(code.n):
                c.reset(i)
        else:
            if self.num_init == 0 and self.init_error:
                for q in range(self.nq):
                    c.id(q, label=IDLE_INIT)
                c.barrier()
            for _ in range(self.num_init):
                for q in range(self.nq):
                    c.reset(q)
            if self.num_init > 0:
                c.barrier()
        if self.basis == "x":
            for i in range(code.n):
                c.h(i)
        if self.initial_state == "-":
            if self.basis == "z":
                for i in code.logical_x[0]:
                    c.x(i)
            else:
                for i in code.logical_z[0]:
                    c.z(i)
        if self.barriers and (self.num_init is None or self.basis == "x"
                              or (self.basis == "z" and self.initial_state == "-")):
            c.barrier()
        start = 0
        for b in range(self.blocks):
            for j, rs in enumerate(self.round_schedule):
                final = b == self.blocks - 1 and j == len(self.round_schedule) - 1
                pauli = self.logical_paulis[j] if self.logical_paulis else ""
                if rs == "x":
                    self.x_round(list(range(start, start + xg)), final, pauli)
                    start += xg
                else:
                    # outcomes, then left flags, then right flags
                    self.z_round(list(range(start, start + zg)), list(range(start + zg, start + 2 * zg)),
                                 list(range(start + 2 * zg, start + 3 * zg)), final, pauli)
                    start += 3 * zg
                if self.logical_paulis and not self.group_meas:
                    self._logical_pauli(self.logical_paulis[j])
                    if self.barriers:
                        c.barrier()
        for i in range(code.n):
            if self.basis == "x" and not self.group_meas:
                c.h(i)
            self._measure(i, start, True)
            start += 1
        return c


def hhc_circuit(d: int = 3, blocks: int = 1, round_schedule: str = "zx", basis: str = "z",
                initial_state: str = "+", barriers: bool = True, idles: bool = True,
                distinct_measurement_idle: bool = True, init_error: bool = True, group_meas: bool = False,
                xprs: bool = False, logical_paulis: str = "", num_initialize: Optional[int] = None,
                idle_before_measure: bool = False) -> EncodedCircuit:
    """``HHCCircuit(HHC(d), ...).syndrome_measurement()`` as an ``EncodedCircuit`` (same options,
    same append order).  Clbits, one register: per round of the schedule the X-gauge outcomes, or
    the Z-gauge outcomes then left flags then right flags; the d*d data-qubit readout last."""
    return _Emitter(d, barriers, idles, distinct_measurement_idle, init_error, group_meas, xprs, blocks,
                    round_schedule, basis, initial_state, logical_paulis, num_initialize,
                    idle_before_measure).build()

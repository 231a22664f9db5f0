"""Synthetic syndrome-extraction gate streams in the ``encoded_circ`` wire format.

These emitters produce the INPUT of the hot path: a
``list[list[int]]`` of opcode rows exactly as the reference's
``CErrorPropagator.load_circuit`` lowers a ``QuantumCircuit``
(reference ``src/qiskit_qec/analysis/cerrorpropagator.py:51-83``), in the order the
reference circuit generators append gates (insertion order is a valid topological
order).  qiskit is not needed.

Followed generator sources (behaviour only, no code shared):
  * SurfaceCodeCircuit   -- reference ``src/qiskit_qec/circuits/surface_code.py:106-141,143-255``
  * RepetitionCodeCircuit-- reference ``src/qiskit_qec/circuits/repetition_code.py:143-217``
  * ArcCircuit           -- reference ``src/qiskit_qec/circuits/repetition_code.py:722-930``
    (explicit links/schedule/color, run_202=False, no conditional resets)
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

# opcode table, reference ``analysis/errorpropagator.h:9-42`` / ``cerrorpropagator.py:21-34``
OP_H, OP_S, OP_X, OP_Y, OP_Z, OP_CX, OP_ID, OP_RESET, OP_MEASURE, OP_BARRIER = range(10)
OP_NAMES = ["h", "s", "x", "y", "z", "cx", "id", "reset", "measure", "barrier"]
NAME_TO_CODE = {n: i for i, n in enumerate(OP_NAMES)}


@dataclass
class EncodedCircuit:
    """A lowered circuit: sizes, opcode rows, and per-row noise labels.

    ``labels[i]`` is what the reference's ``node_name_label`` would return for row i
    (gate label if set else gate name; reference ``utils/dag.py:18-27``).
    ``creg_sizes`` lists classical register sizes in declaration order (used only by
    ``FaultEnumerator.generate()``'s register separators).
    """

    nq: int
    nc: int
    ops: List[List[int]] = field(default_factory=list)
    labels: List[str] = field(default_factory=list)
    creg_sizes: List[int] = field(default_factory=list)
    name: str = ""

    def append(self, name: str, *args: int, label: Optional[str] = None) -> None:
        self.ops.append([NAME_TO_CODE[name], *args])
        self.labels.append(label if label is not None else name)

    # small conveniences mirroring QuantumCircuit method names
    def h(self, q): self.append("h", q)
    def s(self, q): self.append("s", q)
    def x(self, q): self.append("x", q)
    def y(self, q): self.append("y", q)
    def z(self, q): self.append("z", q)
    def id(self, q, label=None): self.append("id", q, label=label)
    def cx(self, c, t): self.append("cx", c, t)
    def reset(self, q): self.append("reset", q)
    def measure(self, q, c): self.append("measure", q, c)
    def barrier(self): self.append("barrier")

    def __len__(self) -> int:
        return len(self.ops)

    def gate_counts(self) -> Dict[str, int]:
        out: Dict[str, int] = {}
        for row in self.ops:
            out[OP_NAMES[row[0]]] = out.get(OP_NAMES[row[0]], 0) + 1
        return out

    def algorithmic_bit_ops(self) -> int:
        """W_bit of SURVEY.md 8(d): 2*n_cx + 2*n_h + n_s + 2*n_reset + n_measure."""
        g = self.gate_counts()
        return (2 * g.get("cx", 0) + 2 * g.get("h", 0) + g.get("s", 0)
                + 2 * g.get("reset", 0) + g.get("measure", 0))


def surface_code_plaquettes(d: int) -> Tuple[List[List[Optional[int]]], List[List[Optional[int]]]]:
    """Plaquette lists (zplaqs, xplaqs) in the reference's scan order.

    Scan y=-1..d-1, x=-1..d-1; keep bulk cells and alternating boundary tabs;
    (x+y) even -> X plaquette with qubit order [p0,p1,p2,p3], otherwise Z plaquette
    with order [p0,p2,p1,p3] (reference ``surface_code.py:106-141``).
    """
    zplaqs: List[List[Optional[int]]] = []
    xplaqs: List[List[Optional[int]]] = []
    for y in range(-1, d):
        for x in range(-1, d):
            in_x = 0 <= x < d - 1
            in_y = 0 <= y < d - 1
            bulk = in_x and in_y
            ztab = (x == -1 and y % 2 == 0) or (x == d - 1 and y % 2 == 1)
            xtab = (y == -1 and x % 2 == 1) or (y == d - 1 and x % 2 == 0)
            if not ((in_x or in_y) and (bulk or ztab or xtab)):
                continue
            corners: List[Optional[int]] = []
            for dy in (0, 1):
                for dx in (0, 1):
                    xx, yy = x + dx, y + dy
                    corners.append(xx + d * yy if (0 <= xx < d and 0 <= yy < d) else None)
            if (x + y) % 2 == 0:
                xplaqs.append([corners[0], corners[1], corners[2], corners[3]])
            else:
                zplaqs.append([corners[0], corners[2], corners[1], corners[3]])
    return zplaqs, xplaqs


def surface_code_circuit(d: int, T: int, basis: str = "z", resets: bool = True,
                         logical: str = "0") -> EncodedCircuit:
    """Gate stream of ``SurfaceCodeCircuit(d, T, basis, resets).circuit[logical]``.

    Qubits: code 0..d^2-1, zplaq next, xplaq last (register declaration order,
    reference ``surface_code.py:78-81,91-93``).  Clbits: per round the zplaq register then
    the xplaq register (``:216-217``); the code_bit register last (``:254``).
    """
    n = d * d
    nxy = (n - 1) // 2
    zplaqs, xplaqs = surface_code_plaquettes(d)
    assert len(zplaqs) == nxy and len(xplaqs) == nxy
    code = list(range(n))
    zq = [n + j for j in range(nxy)]
    xq = [n + nxy + j for j in range(nxy)]
    nc = 2 * nxy * T + (n if T != 0 else 0)
    circ = EncodedCircuit(nq=n + 2 * nxy, nc=nc, name=f"surface_d{d}_T{T}_{basis}_{logical}")
    # preparation (``:143-153``)
    if basis == "z":
        if logical == "1":
            for j in range(d):
                circ.x(code[j * d])
    else:
        for q in code:
            circ.h(q)
        if logical == "1":
            for j in range(d):
                circ.z(code[j])
    for t in range(T):
        final = t == T - 1
        zbits = [2 * nxy * t + j for j in range(nxy)]
        xbits = [2 * nxy * t + nxy + j for j in range(nxy)]
        circ.creg_sizes += [nxy, nxy]
        for q in xq:
            circ.h(q)
        for j in range(4):
            for p, plaq in enumerate(zplaqs):
                if plaq[j] is not None:
                    circ.cx(code[plaq[j]], zq[p])
            for p, plaq in enumerate(xplaqs):
                if plaq[j] is not None:
                    circ.cx(xq[p], code[plaq[j]])
        for q in xq:
            circ.h(q)
        for j in range(nxy):
            circ.measure(xq[j], xbits[j])
            circ.measure(zq[j], zbits[j])
            if resets and not final:
                circ.reset(xq[j])
                circ.reset(zq[j])
    if T != 0:
        if basis == "x":
            for q in code:
                circ.h(q)
        base = 2 * nxy * T
        circ.creg_sizes.append(n)
        for q in code:
            circ.measure(q, base + q)
    return circ


def repetition_code_circuit(d: int, T: int, xbasis: bool = False, resets: bool = False,
                            barriers: bool = False, logical: str = "0") -> EncodedCircuit:
    """Gate stream of ``RepetitionCodeCircuit(d, T, xbasis, resets, barriers=...)``.

    ``QuantumCircuit(link_qubit, code_qubit)`` => link qubits 0..d-2, code qubits d-1..2d-2
    (reference ``repetition_code.py:81-90``).  Delays are not representable (the
    propagator rejects them) and are not emitted.
    """
    link = list(range(d - 1))
    code = [d - 1 + j for j in range(d)]
    nc = (d - 1) * T + (d if T != 0 else 0)
    circ = EncodedCircuit(nq=2 * d - 1, nc=nc, name=f"rep_d{d}_T{T}_{logical}")
    if xbasis:
        for q in code:
            circ.h(q)
    if logical == "1":
        if barriers:
            circ.barrier()
        for q in code:
            (circ.z if xbasis else circ.x)(q)
    for t in range(T):
        final = t == T - 1
        bits = [(d - 1) * t + j for j in range(d - 1)]
        circ.creg_sizes.append(d - 1)
        if barriers:
            circ.barrier()
        if xbasis:
            for q in link:
                circ.h(q)
        for j in range(d - 1):
            if xbasis:
                circ.cx(link[j], code[j])
            else:
                circ.cx(code[j], link[j])
        for j in range(d - 1):
            if xbasis:
                circ.cx(link[j], code[j + 1])
            else:
                circ.cx(code[j + 1], link[j])
        if xbasis:
            for q in link:
                circ.h(q)
        if barriers:
            circ.barrier()
        for j in range(d - 1):
            circ.measure(link[j], bits[j])
        if resets and not final:
            if barriers:
                circ.barrier()
            for j in range(d - 1):
                circ.reset(link[j])
    if T != 0:
        if xbasis:
            for q in code:
                circ.h(q)
        base = (d - 1) * T
        circ.creg_sizes.append(d)
        for j, q in enumerate(code):
            circ.measure(q, base + j)
    return circ


def arc_circuit(links: Sequence[Tuple[int, int, int]], T: int, basis: str = "zx",
                schedule: Optional[Sequence[Sequence[Sequence[int]]]] = None,
                color: Optional[Dict[int, int]] = None, resets: bool = True,
                barriers: bool = True, logical: str = "0") -> EncodedCircuit:
    """Gate stream of ``ArcCircuit(links, T, basis, logical, resets, barriers=..., color=...,
    schedule=..., run_202=False)`` .circuit[basis].

    The reference derives colouring/scheduling with rustworkx matchings
    (``repetition_code.py:621-720``); here both MUST be given explicitly, as the
    reference's own heavy-hex test does (``test/code_circuits/test_rep_codes.py:588-815``).
    Bases containing 'y' emit ``sdg`` (``:733-734``) which the propagator rejects
    (``cerrorpropagator.py:67-68``), so they are refused here.
    Layout (``:746-790``): ``QuantumCircuit(link_qubit, code_qubit)``; code qubits are
    indexed by first appearance in ``links``; clbits: one register of ``len(links)`` bits
    per round, then ``code_bit``.
    """
    if schedule is None or color is None:
        raise ValueError("arc_circuit needs explicit schedule and color (no rustworkx here)")
    if "y" in basis:
        raise ValueError("basis with 'y' needs sdg, which the propagator cannot represent")
    code_qubits: List[int] = []
    link_qubits: List[int] = []
    for c0, a, c1 in links:
        for q in (c0, c1):
            if q not in code_qubits:
                code_qubits.append(q)
        link_qubits.append(a)
    n_code, n_link = len(code_qubits), len(link_qubits)
    code_index = {q: i for i, q in enumerate(code_qubits)}
    link_index = {q: i for i, q in enumerate(link_qubits)}
    code_reg = lambda i: n_link + i  # noqa: E731  (link register is declared first)
    nc = n_link * T + n_code
    circ = EncodedCircuit(nq=n_link + n_code, nc=nc, name=f"arc_{n_code}c{n_link}l_T{T}_{basis}")

    def rotate(c: int, q: int, inverse: bool) -> None:
        if basis[c] == "x":
            circ.h(q)

    def basis_change(inverse: bool) -> None:
        for qubit, i in code_index.items():
            rotate(color[qubit], code_reg(i), inverse)

    if logical == "1":
        for i in range(n_code):
            circ.x(code_reg(i))
    basis_change(False)
    for t in range(T):
        final = t == T - 1
        to_measure = set()
        if barriers:
            circ.barrier()
        for pairs in schedule:
            for qubit_c, qubit_l in pairs:
                qc_, ql_ = code_reg(code_index[qubit_c]), link_index[qubit_l]
                c = color[qubit_c]
                rotate(c, qc_, True)
                circ.cx(qc_, ql_)
                rotate(c, qc_, False)
                to_measure.add(ql_)
        if final:
            basis_change(True)
        if barriers:
            circ.barrier()
        circ.creg_sizes.append(n_link)
        for ql_ in sorted(to_measure):
            circ.measure(ql_, n_link * t + ql_)
        if resets and not final:
            for ql_ in sorted(to_measure):
                circ.reset(ql_)
    if T == 0:
        basis_change(True)
        if barriers:
            circ.barrier()
    circ.creg_sizes.append(n_code)
    for i in range(n_code):
        circ.measure(code_reg(i), n_link * T + i)
    return circ


# 27-qubit heavy-hex (IBM Falcon) coupling map; the ARC uses the degree-2 vertices of
# one bipartition class as link qubits.  The reference supplies no 27-qubit instance
# (its only heavy-hex case is 127 qubits), so this layout is this repo's benchmark input.
HEAVY_HEX_27_EDGES = [
    (0, 1), (1, 2), (1, 4), (2, 3), (3, 5), (4, 7), (5, 8), (6, 7), (7, 10), (8, 9),
    (8, 11), (10, 12), (11, 14), (12, 13), (12, 15), (13, 14), (14, 16), (15, 18),
    (16, 19), (17, 18), (18, 21), (19, 20), (19, 22), (21, 23), (22, 25), (23, 24),
    (24, 25), (25, 26),
]


def heavy_hex_27_arc_layout():
    """(links, schedule, color) for an ARC on the 27-qubit heavy-hex graph.

    Code qubits: 1,3,7,8,12,14,18,19,23,25; link qubits: the degree-2 vertices between
    them (11 links).  schedule = greedy 3-edge-colouring of the (code, link) coupling
    edges in link order; color = BFS 2-colouring of the code-qubit graph from qubit 1.
    """
    nbrs: Dict[int, List[int]] = {}
    for a, b in HEAVY_HEX_27_EDGES:
        nbrs.setdefault(a, []).append(b)
        nbrs.setdefault(b, []).append(a)
    code = {1, 3, 7, 8, 12, 14, 18, 19, 23, 25}
    links = []
    for a in sorted(nbrs):
        if a not in code and len(nbrs[a]) == 2 and all(n in code for n in nbrs[a]):
            c0, c1 = sorted(nbrs[a])
            links.append((c0, a, c1))
    layers: List[List[List[int]]] = []
    busy: List[set] = []
    for c0, a, c1 in links:
        for cq in (c0, c1):
            for k in range(len(layers) + 1):
                if k == len(layers):
                    layers.append([])
                    busy.append(set())
                if cq not in busy[k] and a not in busy[k]:
                    layers[k].append([cq, a])
                    busy[k].update((cq, a))
                    break
    color = {1: 0}
    frontier = [1]
    while frontier:
        nxt = []
        for c in frontier:
            for c0, _, c1 in links:
                for u, v in ((c0, c1), (c1, c0)):
                    if u == c and v not in color:
                        color[v] = 1 - color[c]
                        nxt.append(v)
        frontier = nxt
    return links, layers, color

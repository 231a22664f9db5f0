"""Decoding-graph construction from propagator fault paths.

The reference builds its syndrome graph by running ONE Aer stabilizer simulation per order-1
fault (``decoders/decoding_graph.py:110-163``: ``FaultEnumerator(qc, method="stabilizer")`` ->
outcome string -> ``code.string2nodes`` -> hyperedge).  For the base circuit "0" the noiseless
outcome is all zero, so the stabilizer method's absolute outcomes equal the propagator's flips,
and the whole order-1 enumeration is one GPU launch (``FaultEnumerator.enumerate_packed``).  This
module is the host step that follows it: (outcome bits -> Qiskit result string -> node set ->
hyperedge), restated without qiskit / rustworkx:

  * ``DecodingGraphNode`` / ``DecodingGraphEdge``: the attributes and the equality rule of
    reference ``utils/decoding_graph_attributes.py:26-140``;
  * ``SurfaceCodeNodes`` / ``RepetitionCodeNodes``: ``string2nodes`` of
    ``circuits/surface_code.py:257-415`` and ``circuits/repetition_code.py:218-312`` (pinned by the
    reference's own vectors, ``test/code_circuits/test_surface_codes.py:28-147`` and
    ``test_rep_codes.py:124-160``, in ``tests/test_decoding_cpu.py``);
  * ``SyndromeGraph.add_outcome``: the hyperedge loop of ``decoding_graph.py:131-160``.

Only the distinct outcomes are expanded, in order of first appearance, which gives the node
numbering, edge set and hyperedge list the reference's per-fault loop gives (a repeated outcome
adds nothing there: same nodes, same edges, ``hyperedge not in self.hyperedges`` fails).
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from qiskit_qec_b200.circuits import surface_code_plaquettes


class DecodingGraphNode:
    """Node of a decoding graph (reference ``decoding_graph_attributes.py:26-112``)."""

    def __init__(self, index: int, qubits: Optional[List[int]] = None, is_boundary: bool = False,
                 is_logical: bool = False, time: Optional[int] = None) -> None:
        if not is_boundary and not is_logical and time is None:
            raise ValueError("DecodingGraph node must either have a time or be a boundary or logical node.")
        self.is_boundary = is_boundary
        self.is_logical = is_logical
        self.time = None if (is_boundary or is_logical) else time
        self.qubits = list(qubits) if qubits else []
        self.index = index
        self.properties: Dict[str, Any] = {}

    def _key(self):
        timed = not (self.is_boundary or self.is_logical)
        return (self.index, frozenset(self.qubits), self.is_boundary, self.is_logical,
                self.time if timed else None)

    def __eq__(self, rhs):
        if not isinstance(rhs, DecodingGraphNode):
            return NotImplemented
        return self._key() == rhs._key()

    def __hash__(self) -> int:
        return hash(self._key())

    def __repr__(self) -> str:
        return (f"Node(index={self.index}, qubits={self.qubits}, time={self.time}, "
                f"boundary={self.is_boundary}, logical={self.is_logical})")


@dataclass
class DecodingGraphEdge:
    """Edge of a decoding graph (reference ``decoding_graph_attributes.py:115-140``)."""

    qubits: List[int]
    weight: float
    fault_ids: set = field(default_factory=set)
    properties: Dict[str, Any] = field(default_factory=dict)


def _parity_string(bits: Sequence[int]) -> str:
    return "".join("1" if b else "0" for b in bits)


class SurfaceCodeNodes:
    """``string2nodes`` of the reference ``SurfaceCodeCircuit(d, T, basis, resets)``.

    Result strings are Qiskit's: registers last-declared first, space separated, each with its
    last bit first -- ``"<code_bit> <x_{T-1}> <z_{T-1}> ... <x_0> <z_0>"``
    (``surface_code.py:216-217,254``)."""

    def __init__(self, d: int, T: int, basis: str = "z", resets: bool = True):
        self.d, self.T, self.basis, self.resets = d, T, basis, resets
        self.zplaqs, self.xplaqs = surface_code_plaquettes(d)
        self.x_stabilizer_ops = [[q for q in p if q is not None] for p in self.xplaqs]
        self.z_stabilizer_ops = [[q for q in p if q is not None] for p in self.zplaqs]
        self.logicals = {   # surface_code.py:60-64: x = left / right columns, z = top / bottom rows
            "x": [[j * d for j in range(d)], [(j + 1) * d - 1 for j in range(d)]],
            "z": [list(range(d)), [d * d - 1 - j for j in range(d)]],
        }
        n_plaq = (d * d - 1) // 2
        self.reg_sizes = [n_plaq] * (2 * T) + [d * d]   # declaration order: z_0, x_0, ..., code_bit

    # -- surface_code.py:257-339, on integers: S[r][k] = syndrome of plaquette k in round r
    def _syndromes(self, regs: List[str]) -> List[List[int]]:
        readout = [int(c) for c in regs[0][::-1]]
        plaqs = self.zplaqs if self.basis == "z" else self.xplaqs
        final = [sum(readout[q] for q in p if q is not None) % 2 for p in plaqs]
        measured = regs[2::2] if self.basis == "z" else regs[1::2]   # latest round first
        rounds = [[int(c) for c in reg[::-1]] for reg in measured[::-1]]
        return rounds + [final]

    def _changes(self, S: List[List[int]]) -> List[List[int]]:
        T = self.T
        out = []
        for t in range(len(S)):
            if self.resets:
                row = S[0] if t == 0 else [a ^ b for a, b in zip(S[t], S[t - 1])]
            elif t <= 1:
                row = S[t] if t != T else [a ^ b for a, b in zip(S[t], S[t - 1])]
            elif t == T:
                row = [a ^ b ^ c for a, b, c in zip(S[t], S[t - 1], S[t - 2])]
            else:
                row = [a ^ b for a, b in zip(S[t], S[t - 2])]
            out.append(list(row))
        return out

    def raw_logicals(self, regs: List[str]) -> List[int]:
        readout = [int(c) for c in regs[0][::-1]]
        return [sum(readout[q] for q in self.logicals[self.basis][k]) % 2 for k in (0, 1)]

    def string2nodes(self, string: str, logical: str = "0", all_logicals: bool = False) -> List[DecodingGraphNode]:
        regs = string.split(" ")
        Z = self.raw_logicals(regs)
        nodes = []
        # the reference walks the two logical values last first: index 1 (second logical), then 0
        for k in (1, 0):
            if all_logicals or str(Z[k]) != logical:
                nodes.append(DecodingGraphNode(is_logical=True, is_boundary=True,
                                               qubits=self.logicals[self.basis][k], index=k))
        ops = self.x_stabilizer_ops if self.basis == "x" else self.z_stabilizer_ops
        for t, row in enumerate(self._changes(self._syndromes(regs))):
            for k, flipped in enumerate(row):
                if flipped:
                    nodes.append(DecodingGraphNode(time=t, qubits=ops[k], index=k))
        return nodes


class RepetitionCodeNodes:
    """``string2nodes`` of the reference ``RepetitionCodeCircuit(d, T, resets=...)``
    (``repetition_code.py:218-312``).  Result strings: ``"<code_bit> <link_{T-1}> ... <link_0>"``."""

    def __init__(self, d: int, T: int, resets: bool = False):
        self.d, self.T, self.resets = d, T, resets
        self.gauge_ops = [[j, j + 1] for j in range(d - 1)]
        self.measured_logical = [[0], [d - 1]]
        self.reg_sizes = [d - 1] * T + [d]

    def _rounds(self, string: str) -> List[str]:
        """syndrome strings, latest first: [deduced from the code-bit readout, round T-1, ..., 0]"""
        d = self.d
        measured = string[d:].split(" ")[1:] if len(string) > d else []
        if not self.resets and measured:
            # without resets the link results are cumulative: separate them (:224-234)
            sep = []
            for tt in range(len(measured) - 1):
                sep.append(_parity_string(a != b for a, b in zip(measured[tt], measured[tt + 1])))
            sep.append(measured[-1])
            measured = sep
        final = _parity_string(string[j] != string[j + 1] for j in range(d - 1))
        return [final] + measured

    def string2nodes(self, string: str, logical: str = "0", all_logicals: bool = False) -> List[DecodingGraphNode]:
        d = self.d
        rounds = self._rounds(string)
        nodes = []
        # logical nodes: code qubit 0 is the LAST character of the code-bit register
        for k, ch in enumerate((string[d - 1], string[0])):
            if all_logicals or ch != logical:
                nodes.append(DecodingGraphNode(is_logical=True, is_boundary=True,
                                               qubits=self.measured_logical[k], index=k))
        for t in range(self.T + 1):
            cur = rounds[-t - 1]
            prev = rounds[-t] if t > 0 else "0" * (d - 1)
            for k in range(d - 1):          # string position p holds link d-2-p
                p = d - 2 - k
                if cur[p] != prev[p]:
                    nodes.append(DecodingGraphNode(time=t, qubits=self.gauge_ops[k], index=k))
        return nodes


def outcome_to_string(bits: Sequence[int], reg_sizes: Sequence[int]) -> str:
    """Flat outcome (clbit 0 first) -> Qiskit result string: last clbit first, a space between
    classical registers (what the reference forms from a fault path's outcome,
    ``decoding_graph.py:132``: ``"".join(str(c) for c in output[::-1])``)."""
    parts, j = [], 0
    for size in reg_sizes:
        parts.append("".join("1" if b else "0" for b in bits[j:j + size])[::-1])
        j += size
    return " ".join(parts[::-1])


class SyndromeGraph:
    """Undirected simple graph + hyperedge list, filled outcome by outcome as the reference's
    ``DecodingGraph._make_syndrome_graph`` does (``decoding_graph.py:131-160``).  ``nodes`` is in
    order of first appearance (the reference's node indices), ``edges`` maps an unordered node
    pair to its (latest) edge -- ``rx.PyGraph(multigraph=False)`` semantics."""

    def __init__(self, code):
        self.code = code
        self.nodes: List[DecodingGraphNode] = []
        self._index: Dict[DecodingGraphNode, int] = {}
        self.edges: Dict[Tuple[int, int], DecodingGraphEdge] = {}
        self.hyperedges: List[Dict[Tuple[int, int], DecodingGraphEdge]] = []

    def node_index(self, node: DecodingGraphNode) -> int:
        return self._index[node]

    def add_outcome(self, string: str) -> None:
        nodes = self.code.string2nodes(string)
        for node in nodes:
            if node not in self._index:
                self._index[node] = len(self.nodes)
                self.nodes.append(node)
        hyperedge: Dict[Tuple[int, int], DecodingGraphEdge] = {}
        for source in nodes:
            for target in nodes:
                if target == source:
                    continue
                n0, n1 = self._index[source], self._index[target]
                qubits: List[int] = []
                if not (source.is_logical and target.is_logical):
                    qubits = sorted(set(source.qubits) & set(target.qubits))
                    if not qubits:
                        continue
                if (source.time != target.time and len(qubits) > 1
                        and not source.is_logical and not target.is_logical):
                    qubits = []
                edge = DecodingGraphEdge(qubits, 1)
                self.edges[(min(n0, n1), max(n0, n1))] = edge
                if (n1, n0) not in hyperedge:
                    hyperedge[n0, n1] = edge
        if hyperedge and hyperedge not in self.hyperedges:
            self.hyperedges.append(hyperedge)

    def edge_list(self) -> List[Tuple[int, int]]:
        return sorted(self.edges)


def graph_from_fault_paths(code, fault_paths, reg_sizes: Optional[Sequence[int]] = None) -> SyndromeGraph:
    """The reference's loop, literally: one ``add_outcome`` per fault path.  ``fault_paths`` are
    ``(index, labels, errors, outcome)`` tuples with flat or " "-separated outcomes."""
    reg_sizes = list(reg_sizes or code.reg_sizes)
    graph = SyndromeGraph(code)
    for _, _, _, outcome in fault_paths:
        bits = [c for c in outcome if c != " "]
        graph.add_outcome(outcome_to_string(bits, reg_sizes))
    return graph


def graph_from_packed(code, packed: np.ndarray, n_paths: int,
                      reg_sizes: Optional[Sequence[int]] = None) -> SyndromeGraph:
    """The same graph from the GPU's packed order-1 outcomes (uint32[nc][ceil(n/32)], bit b of
    word w of row c = clbit c of path 32 w + b): distinct outcomes only, in order of first
    appearance."""
    reg_sizes = list(reg_sizes or code.reg_sizes)
    nc = packed.shape[0]
    if sum(reg_sizes) != nc:
        raise ValueError("register sizes do not add up to the number of clbits")
    bits = np.unpackbits(np.ascontiguousarray(packed).view(np.uint8).reshape(nc, -1), axis=1,
                         bitorder="little")[:, :n_paths].T          # [path][clbit]
    _, first = np.unique(np.packbits(bits, axis=1), axis=0, return_index=True)
    graph = SyndromeGraph(code)
    for p in np.sort(first):
        graph.add_outcome(outcome_to_string(bits[p].tolist(), reg_sizes))
    return graph


def decoding_graph_from_enumerator(code, fault_enumerator) -> SyndromeGraph:
    """Syndrome graph of ``code`` from a GPU order-1 ``FaultEnumerator`` of its base circuit "0"
    (one launch for all fault paths, instead of the reference's one Aer run per fault)."""
    if fault_enumerator.order != 1:
        raise ValueError("the syndrome graph is built from order-1 fault paths")
    n = fault_enumerator.num_paths()
    packed, _ = fault_enumerator.enumerate_packed(0, n)
    return graph_from_packed(code, np.asarray(packed), n)

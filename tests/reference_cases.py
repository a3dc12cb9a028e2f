"""The solved trajectories of the reference's documentation (tests/golden/reference_trajectories.npz,
made by tests/golden/make_reference_fixtures.py from /root/reference/docs/data/*.jld2) together with
the quantum systems the documentation builds them for.  Shared by the CPU (oracle) and GPU (CUDA)
tests in tests/test_reference_fixtures.py."""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

X = np.array([[0, 1], [1, 0]], dtype=complex)
Y = np.array([[0, -1j], [1j, 0]], dtype=complex)
Z = np.array([[1, 0], [0, -1]], dtype=complex)
I2 = np.eye(2, dtype=complex)


def qubit(drift=True):
    """docs/literate/man/unitary_problem_templates.jl:22, ket_problem_templates.jl:22,121:
    QuantumSystem(0.1 * PAULIS.Z, [PAULIS.X, PAULIS.Y]) and its driftless twin."""
    return (0.1 * Z if drift else 0 * Z), [X, Y]


def transmon(levels=5, delta=0.2):
    """docs/literate/examples/multilevel_transmon.jl:3-7,47: TransmonSystem(levels=5, delta=0.2) in the
    rotating frame: H = -delta/2 a'a'aa + u1 (a + a') + u2 i(a - a'), everything times 2 pi."""
    a = np.diag(np.sqrt(np.arange(1, levels)), 1).astype(complex)
    ad = a.conj().T
    return 2 * np.pi * (-delta / 2 * ad @ ad @ a @ a), [2 * np.pi * (a + ad), 2 * np.pi * 1j * (a - ad)]


def two_qubit():
    """docs/literate/examples/two_qubit_gates.jl:50-72: J sx sx drift, drives sx_i/2, sy_i/2, times 2 pi."""
    k = np.kron
    return 2 * np.pi * 0.001 * k(X, X), [2 * np.pi * h / 2 for h in (k(X, I2), k(Y, I2), k(I2, X), k(I2, Y))]


# fixture -> (systems, state component names, columns per state, extra)
CASES = {
    "unitary_problem_templates": dict(systems=[qubit()], states=["Ũ⃗"], n=2, unitary=True),
    "unitary_problem_templates_min_time": dict(systems=[qubit()], states=["Ũ⃗"], n=2, unitary=True),
    "ket_problem_templates_state_prob": dict(systems=[qubit()], states=["ψ̃"], n=2, unitary=False),
    "ket_problem_templates_state_prob_min_time": dict(systems=[qubit()], states=["ψ̃"], n=2, unitary=False),
    "ket_problem_templates_sampling_state": dict(systems=[qubit(), qubit(False)],
                                                 states=["ψ̃1_system_1", "ψ̃1_system_2"], n=2, unitary=False),
    "multilevel_transmon_example": dict(systems=[transmon()], states=["Ũ⃗"], n=5, unitary=True),
    "multilevel_transmon_example_leakage": dict(systems=[transmon()], states=["Ũ⃗"], n=5, unitary=True),
    "two_qubit_gates": dict(systems=[two_qubit()], states=["Ũ⃗"], n=4, unitary=True),
    "two_qubit_gates_min_time": dict(systems=[two_qubit()], states=["Ũ⃗"], n=4, unitary=True),
    "two_qubit_gates_molmer": dict(systems=[two_qubit()], states=["Ũ⃗"], n=4, unitary=True),
    "two_qubit_gates_molmer_min_time": dict(systems=[two_qubit()], states=["Ũ⃗"], n=4, unitary=True),
}


def iso(H):
    """G(H) = iso(-iH) = [[Im H, Re H], [-Re H, Im H]]."""
    H = np.asarray(H, dtype=complex)
    return np.block([[H.imag, H.real], [-H.real, H.imag]])


class Case:
    def __init__(self, name):
        g = np.load(os.path.join(GOLD, "reference_trajectories.npz"))
        with open(os.path.join(GOLD, "reference_logs.json")) as fh:
            self.log = json.load(fh).get(name)
        spec = CASES[name]
        self.name = name
        self.data = g[f"{name}.data"]                      # (dim, N)
        self.dim, self.N = self.data.shape
        self.comp = {str(n): (int(a), int(b)) for n, a, b in
                     zip(g[f"{name}.comp_names"], g[f"{name}.comp_start"], g[f"{name}.comp_stop"])}
        self.timestep = str(g[f"{name}.timestep"])
        self.goal = {k.split(".goal.")[1]: g[k] for k in g.files if k.startswith(f"{name}.goal.")}
        self.states = spec["states"]
        self.n = spec["n"]
        self.d = 2 * self.n
        self.c = self.n if spec["unitary"] else 1
        self.unitary = spec["unitary"]
        self.S = len(spec["systems"])
        self.m = len(spec["systems"][0][1])
        self.G_drift = np.stack([iso(h0) for h0, _ in spec["systems"]])
        self.G_drives = np.stack([np.stack([iso(h) for h in hs]) for _, hs in spec["systems"]])
        self.Zvec = np.ascontiguousarray(self.data.T).reshape(-1)   # knot-major flat Z
        self.x_off = [self.comp[s][0] for s in self.states]
        self.u_off = self.comp["u"][0]
        self.dt_off = self.comp[self.timestep][0]

    def rows(self, name):
        a, b = self.comp[name]
        return self.data[a:b]

    def final_operator(self, state=None):
        a, b = self.comp[state or self.states[0]]
        x = self.data[a:b, -1].reshape(self.d, self.c, order="F")
        return x[:self.n] + 1j * x[self.n:]

    def goal_operator(self, state=None):
        x = self.goal[state or self.states[0]].reshape(self.d, self.c, order="F")
        return x[:self.n] + 1j * x[self.n:]

    def reg_terms(self, R=1e-2):
        return [(self.comp[k][0], self.comp[k][1] - self.comp[k][0], R) for k in ("u", "du", "ddu")]

    def derivative_pairs(self):
        return [(self.comp[a][0], self.comp[b][0], self.comp[a][1] - self.comp[a][0]) for a, b in (("u", "du"), ("du", "ddu"))]

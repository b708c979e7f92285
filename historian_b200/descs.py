"""ctypes mirrors of the descriptor structs in include/historian_b200.h, plus
helpers that build them from numpy arrays.

A descriptor only borrows memory: every helper returns (struct, keepalive) and
the caller must keep `keepalive` referenced for as long as the struct is used.
"""
import ctypes as C

import numpy as np

NWEIGHTS = 23
IMM, IMD, IDM, IMI, IIW, EEE = range(6)
STATE_NAMES = ["IMM", "IMD", "IDM", "IMI", "IIW", "EEE"]

COLLAPSE_CHAINS, COUNT_SUBST, COUNT_INDEL, INCLUDE_BEST_TRACE, KEEP_GAPS_OPEN = 1, 2, 4, 8, 16

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_int64_p = C.POINTER(C.c_int64)
c_uint8_p = C.POINTER(C.c_uint8)
c_uint32_p = C.POINTER(C.c_uint32)


class Cell(C.Structure):
    _fields_ = [("xpos", C.c_int32), ("ypos", C.c_int32), ("state", C.c_int32)]


class ProfileDesc(C.Structure):
    _fields_ = [
        ("n_states", C.c_int32), ("n_trans", C.c_int32), ("alph_size", C.c_int32), ("components", C.c_int32),
        ("is_emit", c_uint8_p), ("trans_src", c_int32_p), ("trans_dest", c_int32_p), ("trans_lp", c_double_p),
        ("lp_absorb", c_double_p), ("closest_leaf_pos", c_int32_p),
    ]


class PairHmmDesc(C.Structure):
    _fields_ = [
        ("alph_size", C.c_int32), ("components", C.c_int32), ("lp_trans", C.c_double * NWEIGHTS),
        ("log_root", c_double_p), ("l_sub", c_double_p), ("r_sub", c_double_p),
        ("l_log_ins", c_double_p), ("r_log_ins", c_double_p),
        ("l_log_cpt_weight", c_double_p), ("r_log_cpt_weight", c_double_p),
        ("l_t", C.c_double), ("r_t", C.c_double),
        ("l_ins_wait", C.c_double), ("l_del_wait", C.c_double),
        ("r_ins_wait", C.c_double), ("r_del_wait", C.c_double),
    ]


class EnvDesc(C.Structure):
    _fields_ = [
        ("max_distance", C.c_int32), ("n_cols", C.c_int32), ("cumulative_matches", c_int32_p),
        ("row1_len", C.c_int32), ("row2_len", C.c_int32),
        ("row1_pos_to_col", c_int32_p), ("row2_pos_to_col", c_int32_p),
    ]


class PairInfo(C.Structure):
    _fields_ = [
        ("x_size", C.c_int32), ("y_size", C.c_int32), ("n_cells", C.c_int64), ("n_stored", C.c_int64),
        ("n_diagonals", C.c_int32), ("max_diag_width", C.c_int32), ("h2d_bytes", C.c_int64),
    ]


class ElimResult(C.Structure):
    _fields_ = [
        ("n_states", C.c_int32), ("n_trans", C.c_int32), ("n_via", C.c_int64),
        ("state_cell", C.POINTER(Cell)), ("state_lp_absorb", c_double_p), ("state_fwd_lp", c_double_p),
        ("trans_src", c_int32_p), ("trans_dest", c_int32_p), ("trans_lp", c_double_p),
        ("trans_best_lp", c_double_p), ("trans_best_via_off", c_int64_p), ("trans_best_via", C.POINTER(Cell)),
        ("state_capacity", C.c_int32), ("trans_capacity", C.c_int32), ("via_capacity", C.c_int64),
    ]


class MergeDesc(C.Structure):
    _fields_ = [("node", C.c_int32), ("left", C.c_int32), ("right", C.c_int32),
                ("closest_left", C.c_int32), ("closest_right", C.c_int32), ("closest", C.c_int32),
                ("hmm", C.POINTER(PairHmmDesc))]


class ReconParams(C.Structure):
    _fields_ = [("band", C.c_int32), ("strategy", C.c_int32), ("profile_samples", C.c_int32), ("max_profile_states", C.c_int32),
                ("min_post_prob", C.c_double), ("root_node", C.c_int32), ("mt_state", c_uint32_p)]


RECON_STAGES = ["env", "prepare", "fill", "retry", "trace", "select", "elim", "assemble", "gather", "cleanup"]


class ModelDesc(C.Structure):
    _fields_ = [
        ("alph_size", C.c_int32), ("components", C.c_int32), ("alphabet", C.c_char_p),
        ("cpt_weight", c_double_p), ("ins_prob", c_double_p), ("sub_rate", c_double_p),
        ("eval_re", c_double_p), ("eval_im", c_double_p),
        ("evec_re", c_double_p), ("evec_im", c_double_p),
        ("evec_inv_re", c_double_p), ("evec_inv_im", c_double_p),
        ("branch_sub", c_double_p),
    ]


class TreeDesc(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("parent", c_int32_p), ("branch_length", c_double_p)]


def _arr(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def _ptr(a, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


class Profile:
    """Host-side flattened profile (numpy arrays) — the T1 layout of the C ABI."""

    def __init__(self, is_emit, trans_src, trans_dest, trans_lp, lp_absorb, closest_leaf_pos=None, name=""):
        self.is_emit = _arr(is_emit, np.uint8)
        self.trans_src = _arr(trans_src, np.int32)
        self.trans_dest = _arr(trans_dest, np.int32)
        self.trans_lp = _arr(trans_lp, np.float64)
        self.lp_absorb = _arr(lp_absorb, np.float64)
        assert self.lp_absorb.ndim == 3 and self.lp_absorb.shape[0] == len(self.is_emit)
        self.closest_leaf_pos = None if closest_leaf_pos is None else _arr(closest_leaf_pos, np.int32)
        self.name = name

    @property
    def n_states(self):
        return len(self.is_emit)

    @property
    def n_trans(self):
        return len(self.trans_src)

    @property
    def components(self):
        return self.lp_absorb.shape[1]

    @property
    def alph_size(self):
        return self.lp_absorb.shape[2]

    def desc(self):
        d = ProfileDesc()
        d.n_states, d.n_trans = self.n_states, self.n_trans
        d.alph_size, d.components = self.alph_size, self.components
        d.is_emit = _ptr(self.is_emit, C.c_uint8)
        d.trans_src = _ptr(self.trans_src, C.c_int32)
        d.trans_dest = _ptr(self.trans_dest, C.c_int32)
        d.trans_lp = _ptr(self.trans_lp, C.c_double)
        d.lp_absorb = _ptr(self.lp_absorb, C.c_double)
        if self.closest_leaf_pos is not None:
            d.closest_leaf_pos = _ptr(self.closest_leaf_pos, C.c_int32)
        return d

    @staticmethod
    def from_sequence(seq, alphabet, components=1, wildcard="*"):
        """Leaf profile: Profile(components, alphabet, FastSeq, row), profile.cpp:23-76."""
        n = len(seq) + 2
        A = len(alphabet)
        is_emit = np.ones(n, np.uint8)
        is_emit[0] = is_emit[-1] = 0
        lp = np.full((n, components, A), -np.inf)
        for pos, ch in enumerate(seq):
            tok = -1
            if ch != wildcard:  # tokenize(), fastseq.cpp:10-16: exact char first, then swapped case
                tok = alphabet.find(ch)
                if tok < 0:
                    tok = alphabet.find(ch.swapcase())
            if tok < 0:
                lp[pos + 1, :, :] = 0.0  # wildcard or invalid character (profile.cpp:56-64)
            else:
                lp[pos + 1, :, tok] = 0.0
        src = np.arange(n - 1, dtype=np.int32)
        return Profile(is_emit, src, src + 1, np.zeros(n - 1), lp, np.arange(n, dtype=np.int32).clip(0, len(seq)))


class PairHmm:
    def __init__(self, lp_trans, log_root, l_sub, r_sub, l_log_ins, r_log_ins, l_log_cpt_weight, r_log_cpt_weight,
                 branch=(0, 0, 0, 0, 0, 0)):
        self.lp_trans = _arr(lp_trans, np.float64)
        assert self.lp_trans.shape == (NWEIGHTS,)
        self.log_root = _arr(log_root, np.float64)
        self.l_sub, self.r_sub = _arr(l_sub, np.float64), _arr(r_sub, np.float64)
        self.l_log_ins, self.r_log_ins = _arr(l_log_ins, np.float64), _arr(r_log_ins, np.float64)
        self.l_log_cpt_weight = _arr(l_log_cpt_weight, np.float64)
        self.r_log_cpt_weight = _arr(r_log_cpt_weight, np.float64)
        self.branch = tuple(float(b) for b in branch)

    @property
    def components(self):
        return self.log_root.shape[0]

    @property
    def alph_size(self):
        return self.log_root.shape[1]

    def desc(self):
        d = PairHmmDesc()
        d.alph_size, d.components = self.alph_size, self.components
        for k in range(NWEIGHTS):
            d.lp_trans[k] = self.lp_trans[k]
        d.log_root = _ptr(self.log_root, C.c_double)
        d.l_sub, d.r_sub = _ptr(self.l_sub, C.c_double), _ptr(self.r_sub, C.c_double)
        d.l_log_ins, d.r_log_ins = _ptr(self.l_log_ins, C.c_double), _ptr(self.r_log_ins, C.c_double)
        d.l_log_cpt_weight = _ptr(self.l_log_cpt_weight, C.c_double)
        d.r_log_cpt_weight = _ptr(self.r_log_cpt_weight, C.c_double)
        d.l_t, d.r_t, d.l_ins_wait, d.l_del_wait, d.r_ins_wait, d.r_del_wait = self.branch
        return d

    @staticmethod
    def from_probs(l, r, root, cpt_weight):
        """PairHMM(l, r, root), pairhmm.cpp:5-44.  l, r: dicts with ins, del, insExt, delExt,
        sub [C][A][A], insProb [C][A], t, insWait, delWait.  root: [C][A] probabilities."""
        lI, lD, lIx, lDx = l["ins"], l["del"], l["insExt"], l["delExt"]
        rI, rD, rIx, rDx = r["ins"], r["del"], r["insExt"], r["delExt"]
        with np.errstate(divide="ignore"):
            lg = lambda v: float(np.log(v))
            w = np.zeros(NWEIGHTS)
            w[0] = lg((1 - lI) * (1 - rI) * (1 - lD) * (1 - rD))   # imm_imm
            w[1] = lg((1 - lI) * (1 - rI) * (1 - lD) * rD)         # imm_imd
            w[2] = lg((1 - lI) * (1 - rI) * lD * (1 - rD))         # imm_idm
            w[3] = lg(rI)                                           # imm_imi
            w[4] = lg(lI * (1 - rI))                                # imm_iiw
            w[5] = lg((1 - lI) * (1 - rI))                          # imm_eee
            w[6] = lg((1 - lI) * (1 - lD) * (1 - rDx))             # imd_imm
            w[7] = lg((1 - lI) * (1 - lD) * rDx)                   # imd_imd
            w[8] = lg((1 - lI) * lD * (1 - rDx))                   # imd_idm
            w[9] = lg((1 - lI) * (1 - rDx))                        # imd_eee
            w[10] = lg((1 - rI) * (1 - lDx) * (1 - rD))            # idm_imm
            w[11] = lg((1 - rI) * (1 - lDx) * rD)                  # idm_imd
            w[12] = lg((1 - rI) * lDx * (1 - rD))                  # idm_idm
            w[13] = lg((1 - rI) * (1 - lDx))                       # idm_eee
            w[14] = lg((1 - lI) * (1 - rIx) * (1 - lD) * (1 - rD)) # imi_imm
            w[15] = lg((1 - lI) * (1 - rIx) * (1 - lD) * rD)       # imi_imd
            w[16] = lg(rIx)                                         # imi_imi
            w[17] = lg(lI * (1 - rIx))                              # imi_iiw
            w[18] = lg((1 - lI) * (1 - rIx))                        # imi_eee
            w[19] = lg((1 - lIx) * (1 - lD) * (1 - rD))            # iiw_imm
            w[20] = lg((1 - lIx) * lD * (1 - rD))                  # iiw_idm
            w[21] = lg(lIx)                                         # iiw_iiw
            w[22] = lg(1 - lIx)                                     # iiw_eee
            log_cpt = np.log(np.asarray(cpt_weight, float))
            log_root = np.log(np.asarray(root, float)) + log_cpt[:, None]
            return PairHmm(w, log_root, l["sub"], r["sub"], np.log(np.asarray(l["insProb"], float)),
                           np.log(np.asarray(r["insProb"], float)), log_cpt, log_cpt,
                           (l.get("t", 0), r.get("t", 0), l.get("insWait", 0), l.get("delWait", 0),
                            r.get("insWait", 0), r.get("delWait", 0)))


class Envelope:
    def __init__(self, max_distance=-1, cumulative_matches=None, row1_pos_to_col=None, row2_pos_to_col=None):
        self.max_distance = int(max_distance)
        self.cum = None if cumulative_matches is None else _arr(cumulative_matches, np.int32)
        self.r1 = None if row1_pos_to_col is None else _arr(row1_pos_to_col, np.int32)
        self.r2 = None if row2_pos_to_col is None else _arr(row2_pos_to_col, np.int32)

    def desc(self):
        d = EnvDesc()
        d.max_distance = self.max_distance
        if self.max_distance >= 0:
            d.n_cols = len(self.cum) - 1
            d.cumulative_matches = _ptr(self.cum, C.c_int32)
            d.row1_len, d.row2_len = len(self.r1) - 1, len(self.r2) - 1
            d.row1_pos_to_col = _ptr(self.r1, C.c_int32)
            d.row2_pos_to_col = _ptr(self.r2, C.c_int32)
        return d

    @staticmethod
    def from_guide_rows(row1, row2, max_distance):
        """GuideAlignmentEnvelope(guide,row1,row2,maxDistance), alignpath.cpp:282-310.
        row1/row2: boolean arrays over guide columns (True = residue)."""
        row1, row2 = np.asarray(row1, bool), np.asarray(row2, bool)
        cum = np.concatenate([[0], np.cumsum(row1 & row2)]).astype(np.int32)
        r1 = np.concatenate([[0], np.nonzero(row1)[0] + 1]).astype(np.int32)
        r2 = np.concatenate([[0], np.nonzero(row2)[0] + 1]).astype(np.int32)
        return Envelope(max_distance, cum, r1, r2)


class QuickAlignScores(C.Structure):
    """hist_quickalign_scores (quickalign.cpp:22-50)."""
    _fields_ = [("alph_size", C.c_int32), ("submat", C.POINTER(C.c_double))] + [(k, C.c_double) for k in (
        "no_gap", "gap_open", "gap_extend", "m2m", "m2i", "m2d", "i2i", "i2m", "i2d", "d2d", "d2m")]


class QuickAlignPair(C.Structure):
    """hist_quickalign_pair."""
    _fields_ = [("x_len", C.c_int32), ("y_len", C.c_int32), ("x_tok", C.POINTER(C.c_int8)), ("y_tok", C.POINTER(C.c_int8)),
                ("n_diagonals", C.c_int32), ("diagonals", C.POINTER(C.c_int32))]


class BranchModel(C.Structure):
    """hist_branch_model."""
    _fields_ = [("alph_size", C.c_int32), ("components", C.c_int32), ("log_sub", C.POINTER(C.c_double)), ("log_ins", C.POINTER(C.c_double)),
                ("log_cpt_weight", C.POINTER(C.c_double)), ("lp_trans", (C.c_double * 4) * 3)]


class GuideParams(C.Structure):
    """hist_guide_params (DiagEnvParams, diagenv.h:24-35, plus AlignGraph's choice of edges and the generator's seed)."""
    _fields_ = [("all_pairs", C.c_int32), ("seed", C.c_uint32), ("sparse", C.c_int32), ("kmer_len", C.c_int32), ("kmer_threshold", C.c_int32),
                ("band_size", C.c_int32), ("cell_size", C.c_int32), ("reserved", C.c_int32), ("max_size", C.c_int64)]


def read_dump(path):
    """Reads the binary container written by oracle/ref_dump.cpp into {name: ndarray|str}."""
    out = {}
    dt = {"d": np.float64, "i": np.int32, "b": np.uint8, "u": np.uint32}
    with open(path, "rb") as f:
        data = f.read()
    off = 0
    while off < len(data):
        nl = int(np.frombuffer(data, np.uint32, 1, off)[0]); off += 4
        name = data[off:off + nl].decode(); off += nl
        ty = chr(data[off]); off += 1
        nd = int(np.frombuffer(data, np.uint32, 1, off)[0]); off += 4
        shape = tuple(int(v) for v in np.frombuffer(data, np.uint64, nd, off)); off += 8 * nd
        nb = int(np.frombuffer(data, np.uint64, 1, off)[0]); off += 8
        raw = data[off:off + nb]; off += nb
        if ty == "s":
            out[name] = raw.decode()
        else:
            out[name] = np.frombuffer(raw, dt[ty]).reshape(shape).copy()
    return out


def profile_from_dump(d, pfx):
    return Profile(d[pfx + "is_emit"], d[pfx + "trans_src"], d[pfx + "trans_dest"], d[pfx + "trans_lp"],
                   d[pfx + "lp_absorb"], d.get(pfx + "closest_leaf_pos"), name=str(d.get(pfx + "name", "")))


def hmm_from_dump(d, pfx):
    return PairHmm(d[pfx + "lp_trans"], d[pfx + "log_root"], d[pfx + "l_sub"], d[pfx + "r_sub"], d[pfx + "l_log_ins"],
                   d[pfx + "r_log_ins"], d[pfx + "l_log_cpt_weight"], d[pfx + "r_log_cpt_weight"], d[pfx + "branch"])


def env_from_dump(d, pfx):
    md = int(d[pfx + "max_distance"][0])
    if md < 0:
        return Envelope(-1)
    return Envelope(md, d[pfx + "cumulative_matches"], d[pfx + "row1_pos_to_col"], d[pfx + "row2_pos_to_col"])


class SumProdModel:
    """hist_model_desc: a RateModel in eigen-decomposed form (EigenModel, model.cpp:1135-1199)."""

    def __init__(self, alphabet, cpt_weight, ins_prob, sub_rate, evals, evec, evec_inv, branch_sub=None):
        self.alphabet = alphabet if isinstance(alphabet, bytes) else alphabet.encode()
        self.cpt_weight = _arr(cpt_weight, np.float64)
        self.ins_prob = _arr(ins_prob, np.float64)
        self.sub_rate = _arr(sub_rate, np.float64)
        ev, v, vi = np.asarray(evals, complex), np.asarray(evec, complex), np.asarray(evec_inv, complex)
        self.parts = [_arr(a, np.float64) for a in (ev.real, ev.imag, v.real, v.imag, vi.real, vi.imag)]
        self.branch_sub = None if branch_sub is None else _arr(branch_sub, np.float64)

    @property
    def components(self):
        return self.ins_prob.shape[0]

    @property
    def alph_size(self):
        return self.ins_prob.shape[1]

    def desc(self):
        d = ModelDesc()
        d.alph_size, d.components = self.alph_size, self.components
        d.alphabet = self.alphabet
        d.cpt_weight, d.ins_prob, d.sub_rate = (_ptr(a, C.c_double) for a in (self.cpt_weight, self.ins_prob, self.sub_rate))
        (d.eval_re, d.eval_im, d.evec_re, d.evec_im, d.evec_inv_re, d.evec_inv_im) = (_ptr(a, C.c_double) for a in self.parts)
        if self.branch_sub is not None:
            d.branch_sub = _ptr(self.branch_sub, C.c_double)
        return d

    @staticmethod
    def from_golden(d, use_branch_sub=True):
        return SumProdModel(bytes(d["alphabet"]), d["cpt_weight"], d["ins_prob"], d["sub_rate"], d["eval_re"] + 1j * d["eval_im"],
                            d["evec_re"] + 1j * d["evec_im"], d["evec_inv_re"] + 1j * d["evec_inv_im"],
                            d["branch_sub"] if use_branch_sub else None)

    @staticmethod
    def from_rate_model(model):
        """Eigen-decompose a model.RateModel with numpy (host-side model preparation, O(C A^3) once)."""
        from .model import EigenModel
        e = EigenModel(model)
        return SumProdModel(model.alphabet, model.cptWeight, model.insProb, model.subRate, np.array(e.eval), np.array(e.evec),
                            np.array(e.evecInv))


class TreeTopology:
    """hist_tree_desc: nodes in postorder, children < parent, root last (tree.cpp:141)."""

    def __init__(self, parent, branch_length):
        self.parent = _arr(parent, np.int32)
        self.branch_length = _arr(branch_length, np.float64)

    @property
    def n_nodes(self):
        return len(self.parent)

    def desc(self):
        d = TreeDesc()
        d.n_nodes = self.n_nodes
        d.parent, d.branch_length = _ptr(self.parent, C.c_int32), _ptr(self.branch_length, C.c_double)
        return d


def counts_size(components, alph_size):
    """Doubles in the flat EigenCounts buffer: 7 indel slots + rootCount[C][A] + eigenCount[C][A][A] (re, im)."""
    return 7 + components * alph_size + 2 * components * alph_size * alph_size

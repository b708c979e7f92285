"""ctypes binding of libhistorian_b200.so (the C ABI in include/historian_b200.h).

There is no CPU fallback: loading fails loudly if the CUDA library has not been
built, and Context() raises if no CUDA device is usable.
"""
import ctypes as C
import os

import numpy as np

from . import descs as D

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HIST_B200_LIB") or os.path.join(_HERE, "libhistorian_b200.so")   # override: kernel tuning variants
_lib = None


class HistError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("historian_b200 error %d: %s" % (code, msg))
        self.code = code


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s not found: build it with `python -m historian_b200.build` (needs nvcc). "
                          "There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.hist_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.hist_ctx_destroy.argtypes = [vp]
    lib.hist_ctx_destroy.restype = None
    lib.hist_last_error.argtypes = [vp]
    lib.hist_last_error.restype = C.c_char_p
    lib.hist_ctx_stream.argtypes = [vp]
    lib.hist_ctx_stream.restype = vp
    lib.hist_ctx_synchronize.argtypes = [vp]
    lib.hist_ctx_launch_count.argtypes = [vp]
    lib.hist_ctx_launch_count.restype = C.c_int64
    lib.hist_quickalign_batch.argtypes = [vp, C.c_int32, C.POINTER(D.QuickAlignPair), C.POINTER(D.QuickAlignScores), D.c_double_p,
                                          D.c_int32_p, D.c_int32_p, D.c_uint8_p, D.c_int64_p, D.c_int32_p]
    lib.hist_distance_matrix.argtypes = [vp, C.POINTER(D.ModelDesc), C.c_double, C.c_int32, C.c_int32, C.c_char_p, C.c_int32, D.c_double_p]
    lib.hist_build_tree.argtypes = [vp, C.c_int32, D.c_double_p, C.c_int32, C.c_double, D.c_int32_p, D.c_double_p]
    lib.hist_guide_align.argtypes = [vp, C.c_int32, C.POINTER(C.c_int8), D.c_int64_p, C.POINTER(D.QuickAlignScores), C.POINTER(D.GuideParams),
                                     D.c_int32_p, D.c_int32_p, D.c_int32_p, D.c_int32_p, D.c_double_p, C.c_int64, D.c_int32_p]
    lib.hist_align_path_merge.argtypes = [vp, C.c_int32, D.c_int32_p, D.c_int32_p, D.c_int32_p, C.POINTER(C.c_uint8), C.c_int32, D.c_int64_p,
                                          D.c_int32_p, D.c_int32_p]
    lib.hist_sumprod_excluded_posteriors.argtypes = [vp, C.c_char_p, C.c_int64, C.c_int32, D.c_int32_p, D.c_int32_p, C.c_int32, D.c_double_p]
    lib.hist_branch_align.argtypes = [vp, C.POINTER(D.BranchModel), C.c_int32, D.c_double_p, C.c_int32, D.c_double_p, C.c_int32, D.c_int32_p,
                                      D.c_int32_p, D.c_double_p, C.POINTER(C.c_uint8), D.c_int32_p]
    lib.hist_diagonal_envelope.argtypes = [vp, C.POINTER(C.c_int8), C.c_int32, C.POINTER(C.c_int8), C.c_int32, C.c_int32, C.POINTER(D.GuideParams),
                                           D.c_int32_p, C.c_int32, D.c_int32_p]
    lib.hist_ctx_h2d_bytes.argtypes = [vp]
    lib.hist_ctx_h2d_bytes.restype = C.c_int64
    lib.hist_ctx_set_timing.argtypes = [vp, C.c_int]
    lib.hist_ctx_fill_time_ms.argtypes = [vp, D.c_double_p, D.c_int64_p]
    lib.hist_profile_upload.argtypes = [vp, C.POINTER(D.ProfileDesc), C.POINTER(vp)]
    lib.hist_pairs_destroy.argtypes = [C.POINTER(vp), C.c_int32]
    lib.hist_pairs_destroy.restype = None
    lib.hist_profiles_destroy.argtypes = [C.POINTER(vp), C.c_int32]
    lib.hist_profiles_destroy.restype = None
    lib.hist_profile_from_residues.argtypes = [vp, C.c_int32, C.POINTER(C.c_char_p), D.c_int32_p, C.c_char_p, C.c_int32, C.POINTER(vp)]
    lib.hist_profile_from_sequences.argtypes = [vp, C.c_int32, C.POINTER(vp), D.c_int32_p, C.c_int32, C.c_int32, C.POINTER(vp)]
    lib.hist_profile_sum_path_absorb.argtypes = [vp, D.c_double_p, D.c_double_p, D.c_double_p, D.c_double_p]
    lib.hist_profile_sum_path_absorb_noalloc.argtypes = [vp, D.c_double_p, D.c_double_p, D.c_double_p, D.c_double_p]
    lib.hist_noalloc_scratch_bytes.argtypes = []
    lib.hist_noalloc_scratch_bytes.restype = C.c_int64
    lib.hist_profile_destroy.argtypes = [vp]
    lib.hist_profile_destroy.restype = None
    lib.hist_pair_prepare.argtypes = [vp, vp, vp, C.POINTER(D.PairHmmDesc), C.POINTER(D.EnvDesc), C.POINTER(vp)]
    lib.hist_pair_prepare_batch.argtypes = [vp, C.c_int32, C.POINTER(vp), C.POINTER(vp), C.POINTER(C.POINTER(D.PairHmmDesc)),
                                            C.POINTER(C.POINTER(D.EnvDesc)), C.c_int32, C.POINTER(vp)]
    lib.hist_pair_destroy.argtypes = [vp]
    lib.hist_pair_destroy.restype = None
    lib.hist_pair_get_info.argtypes = [vp, C.POINTER(D.PairInfo)]
    lib.hist_pair_fetch_tables.argtypes = [vp] + [D.c_double_p] * 6
    lib.hist_pair_fetch_envelope.argtypes = [vp, D.c_int32_p, D.c_int32_p, D.c_uint8_p, D.c_uint8_p, D.c_uint8_p]
    lib.hist_fwd_fill.argtypes = [vp, D.c_double_p]
    lib.hist_bwd_fill.argtypes = [vp, D.c_double_p]
    lib.hist_fill_batch.argtypes = [vp, C.POINTER(vp), C.c_int32, C.c_int, D.c_double_p, D.c_double_p]
    lib.hist_fill_batch_results.argtypes = [vp, C.POINTER(vp), C.c_int32, C.c_int, D.c_double_p, D.c_double_p]
    lib.hist_cells_fetch.argtypes = [vp, C.c_int, D.c_double_p, D.c_int32_p, D.c_int32_p, C.c_int64, D.c_int64_p]
    lib.hist_trace_best.argtypes = [vp, D.Cell, C.POINTER(D.Cell), C.c_int64, D.c_int64_p]
    lib.hist_trace_best_forward.argtypes = [vp, D.Cell, C.POINTER(D.Cell), C.c_int64, D.c_int64_p]
    lib.hist_trace_sample.argtypes = [vp, D.c_uint32_p, C.c_int32, C.POINTER(D.Cell), C.c_int64, D.c_int64_p]
    lib.hist_posterior_cells.argtypes = [vp, C.c_double, C.POINTER(D.Cell), D.c_double_p, C.c_int64, D.c_int64_p]
    lib.hist_fb_counts.argtypes = [vp, D.c_double_p, D.c_double_p, D.c_double_p, D.c_double_p, C.POINTER(D.Cell), D.c_double_p,
                                   C.c_int64, D.c_int64_p]
    lib.hist_profile_eliminate.argtypes = [vp, C.POINTER(D.Cell), C.c_int64, C.c_int32, C.POINTER(D.ElimResult)]
    lib.hist_recon_create.argtypes = [vp, C.c_int32, C.c_char_p, C.c_int32, C.POINTER(vp)]
    lib.hist_recon_destroy.argtypes = [vp]
    lib.hist_recon_destroy.restype = None
    lib.hist_recon_set_guide.argtypes = [vp, C.c_int32, C.c_int32, D.c_int32_p, D.c_uint8_p]
    lib.hist_recon_add_leaves.argtypes = [vp, C.c_int32, D.c_int32_p, C.POINTER(C.c_char_p), D.c_int32_p]
    lib.hist_recon_level.argtypes = [vp, C.c_int32, C.POINTER(D.MergeDesc), C.POINTER(D.ReconParams), D.c_double_p]
    lib.hist_recon_alignment.argtypes = [vp, D.c_int32_p, D.c_uint8_p, C.c_int64]
    lib.hist_recon_stage_ms.argtypes = [vp, D.c_double_p, C.c_int]
    lib.hist_recon_profile.argtypes = [vp, C.c_int32]
    lib.hist_recon_profile.restype = vp
    lib.hist_recon_export_node.argtypes = [vp, C.c_int32, vp, C.c_int64, D.c_int64_p, vp, D.c_int64_p]
    lib.hist_recon_import_node.argtypes = [vp, C.c_int32, vp, C.c_int64, vp]
    lib.hist_recon_release_node.argtypes = [vp, C.c_int32]
    lib.hist_recon_export_stores.argtypes = [vp, vp, C.c_int64, D.c_int64_p]
    lib.hist_recon_import_stores.argtypes = [vp, vp, C.c_int64]
    lib.hist_profile_get_sizes.argtypes = [vp, D.c_int32_p, D.c_int32_p, D.c_int32_p, D.c_int32_p]
    lib.hist_profile_fetch.argtypes = [vp, D.c_uint8_p, D.c_int32_p, D.c_int32_p, D.c_double_p, D.c_double_p, D.c_int32_p]
    lib.hist_sumprod_create.argtypes = [vp, C.POINTER(D.ModelDesc), C.POINTER(D.TreeDesc), C.POINTER(vp)]
    lib.hist_sumprod_destroy.argtypes = [vp]
    lib.hist_sumprod_destroy.restype = None
    lib.hist_sumprod_counts_size.argtypes = [vp]
    lib.hist_sumprod_counts_size.restype = C.c_size_t
    lib.hist_sumprod_columns.argtypes = [vp, vp, vp, C.c_int64, C.c_int32, vp, vp, vp, vp]
    lib.hist_sumprod_debug_column.argtypes = [vp, C.c_char_p] + [D.c_double_p] * 7
    lib.hist_sumprod_transform_counts.argtypes = [vp, vp, D.c_double_p, D.c_double_p]
    lib.hist_sumprod_branch_tables.argtypes = [vp, D.c_double_p, D.c_double_p]
    _lib = lib
    return lib


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


class Context:
    def __init__(self, device=0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.hist_ctx_create(device, C.byref(h))
        if rc != 0:
            raise HistError(rc, self.lib.hist_last_error(None).decode())
        self.h = h
        self.device = device

    def check(self, rc):
        if rc != 0:
            raise HistError(rc, self.lib.hist_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.lib.hist_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.lib.hist_ctx_stream(self.h)

    def synchronize(self):
        self.check(self.lib.hist_ctx_synchronize(self.h))

    @property
    def launch_count(self):
        return int(self.lib.hist_ctx_launch_count(self.h))

    def quickalign(self, pairs, submat, scores, want_paths=True):
        """QuickAlignMatrix (quickalign.cpp:7-214) for a batch of pairs: hist_quickalign_batch.
        pairs: [(x_tok int8[], y_tok int8[], diagonals int32[])]; submat [A][A]; scores: the 11 scalars
        (noGap, gapOpen, gapExtend, m2m, m2i, m2d, i2i, i2m, i2d, d2d, d2m).
        Returns (result[n], ends[n][2], [path uint8[] per pair] or None)."""
        n = len(pairs)
        sub = np.ascontiguousarray(submat, np.float64)
        sc = D.QuickAlignScores(int(sub.shape[0]), _p(sub, C.c_double), *[float(v) for v in scores])
        keep, arr = [], (D.QuickAlignPair * n)()
        off = np.zeros(n + 1, np.int64)
        for k, (xt, yt, dg) in enumerate(pairs):   # dg None: the full envelope
            xt, yt = np.ascontiguousarray(xt, np.int8), np.ascontiguousarray(yt, np.int8)
            dg = None if dg is None else np.ascontiguousarray(dg, np.int32)
            keep.append((xt, yt, dg))
            arr[k] = D.QuickAlignPair(len(xt), len(yt), _p(xt, C.c_int8), _p(yt, C.c_int8), len(xt) + len(yt) - 1 if dg is None else len(dg), _p(dg, C.c_int32))
            off[k + 1] = off[k] + len(xt) + len(yt)
        result, xe, ye = np.zeros(n), np.zeros(n, np.int32), np.zeros(n, np.int32)
        buf = np.zeros(max(1, int(off[n])), np.uint8) if want_paths else None
        plen = np.zeros(n, np.int32)
        self.check(self.lib.hist_quickalign_batch(self.h, n, arr, C.byref(sc), _p(result, C.c_double), _p(xe, C.c_int32), _p(ye, C.c_int32),
                                                  _p(buf, C.c_uint8) if want_paths else None, _p(off, C.c_int64) if want_paths else None,
                                                  _p(plen, C.c_int32) if want_paths else None))
        paths = [buf[off[k]:off[k] + plen[k]].copy() for k in range(n)] if want_paths else None
        return result, np.stack([xe, ye], 1), paths

    def guide_align(self, seq_tokens, submat, scores, all_pairs=False, seed=5489, sparse=True, kmer_len=6, kmer_threshold=-1, band_size=64,
                    cell_size=40, max_size=None):
        """AlignGraph::mstAlign (span.cpp:35-158) of unaligned sequences: hist_guide_align.  seq_tokens: int8 token arrays
        (< 0: not in the alphabet); submat / scores as for quickalign; the rest is hist_guide_params (defaults: the reference's
        random spanning graph, generator seed and DiagEnvParams; max_size None = this machine's memory, DiagEnvParams::effectiveMaxSize,
        diagenv.cpp:79-89).  Returns (bits uint8 [n][cols], info) with info = {"edges": [E][2], "edge_lp": [E], "mst": [n-1]}."""
        toks = [np.ascontiguousarray(t, np.int8) for t in seq_tokens]
        n = len(toks)
        off = np.zeros(n + 1, np.int64)
        off[1:] = np.cumsum([len(t) for t in toks])
        flat = np.ascontiguousarray(np.concatenate(toks), np.int8)
        sub = np.ascontiguousarray(submat, np.float64)
        sc = D.QuickAlignScores(int(sub.shape[0]), _p(sub, C.c_double), *[float(v) for v in scores])
        if max_size is None:
            max_size = os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES")
        gp = D.GuideParams(int(bool(all_pairs)), int(seed), int(bool(sparse)), int(kmer_len), int(kmer_threshold), int(band_size), int(cell_size), 0, int(max_size))
        cap = n * (n - 1) // 2
        rcol, ncols, nedges = np.zeros(int(off[n]), np.int32), np.zeros(1, np.int32), np.zeros(1, np.int32)
        erows, elp, mst = np.zeros((max(1, cap), 2), np.int32), np.zeros(max(1, cap)), np.zeros(max(1, n - 1), np.int32)
        self.check(self.lib.hist_guide_align(self.h, n, _p(flat, C.c_int8), _p(off, C.c_int64), C.byref(sc), C.byref(gp), _p(rcol, C.c_int32),
                                             _p(ncols, C.c_int32), _p(nedges, C.c_int32), _p(erows, C.c_int32), _p(elp, C.c_double), cap, _p(mst, C.c_int32)))
        bits = np.zeros((n, int(ncols[0])), np.uint8)
        for s in range(n):
            bits[s, rcol[off[s]:off[s + 1]]] = 1
        e = int(nedges[0])
        return bits, {"edges": erows[:e].copy(), "edge_lp": elp[:e].copy(), "mst": mst[:max(0, n - 1)].copy()}

    def branch_align(self, log_sub, log_ins, log_cpt_weight, lp_trans, x_pwm, y_pwm, max_distance=-1, x_key=None, y_key=None):
        """Refiner::BranchMatrix + best() (refiner.cpp:10-106): hist_branch_align.  log_sub [C][A][A], log_ins [C][A], log_cpt_weight [C],
        lp_trans [3][4]; x_pwm [x_len][C][A], y_pwm [y_len][C][A].  Returns (lp_end, path uint8[]: 3 both, 1 parent only, 2 child only)."""
        log_sub, log_ins, lcw = (np.ascontiguousarray(v, np.float64) for v in (log_sub, log_ins, log_cpt_weight))
        Cc, A = log_ins.shape
        bm = D.BranchModel(A, Cc, _p(log_sub, C.c_double), _p(log_ins, C.c_double), _p(lcw, C.c_double))
        for a in range(3):
            for b in range(4):
                bm.lp_trans[a][b] = float(lp_trans[a][b])
        x_pwm, y_pwm = np.ascontiguousarray(x_pwm, np.float64).reshape(-1, Cc, A), np.ascontiguousarray(y_pwm, np.float64).reshape(-1, Cc, A)
        xl, yl = len(x_pwm), len(y_pwm)
        kx = None if x_key is None else np.ascontiguousarray(x_key, np.int32)
        ky = None if y_key is None else np.ascontiguousarray(y_key, np.int32)
        assert max_distance < 0 or (len(kx) == xl + 1 and len(ky) == yl + 1)
        lp, path, plen = np.zeros(1), np.zeros(max(1, xl + yl), np.uint8), np.zeros(1, np.int32)
        self.check(self.lib.hist_branch_align(self.h, C.byref(bm), xl, _p(x_pwm, C.c_double), yl, _p(y_pwm, C.c_double), int(max_distance),
                                              _p(kx, C.c_int32), _p(ky, C.c_int32), _p(lp, C.c_double), _p(path, C.c_uint8), _p(plen, C.c_int32)))
        return float(lp[0]), path[:int(plen[0])].copy()

    def distance_matrix(self, sp_model, expected_sub_rate, rows, max_iterations=100):
        """RateModel::distanceMatrix (model.cpp:533-553): hist_distance_matrix.  sp_model: descs.SumProdModel (alphabet, component
        weights, eigen-decomposition); rows: uint8 [n][cols] gapped characters.  Returns dist [n][n]."""
        rows = np.ascontiguousarray(rows, np.uint8)
        n, cols = rows.shape
        dist = np.zeros((n, n))
        md = sp_model.desc()
        self.check(self.lib.hist_distance_matrix(self.h, C.byref(md), float(expected_sub_rate), n, cols, rows.ctypes.data_as(C.c_char_p),
                                                 int(max_iterations), _p(dist, C.c_double)))
        return dist

    def build_tree(self, dist, method=0, min_branch_length=1e-9):
        """Tree::buildByUPGMA (method 0) / buildByNeighborJoining (method 1): hist_build_tree.  Returns (child [2n-1][2], branch_length [2n-1])."""
        dist = np.ascontiguousarray(dist, np.float64)
        n = len(dist)
        child, blen = np.zeros((2 * n - 1, 2), np.int32), np.zeros(2 * n - 1)
        self.check(self.lib.hist_build_tree(self.h, n, _p(dist, C.c_double), int(method), float(min_branch_length), _p(child, C.c_int32), _p(blen, C.c_double)))
        return child, blen

    @property
    def h2d_bytes(self):
        """Bytes copied host -> device so far for profiles and DPMatrix constructors (hist_ctx_h2d_bytes)."""
        return int(self.lib.hist_ctx_h2d_bytes(self.h))

    def set_timing(self, enabled=True):
        self.check(self.lib.hist_ctx_set_timing(self.h, int(enabled)))

    def fill_time_ms(self):
        """(summed fill-kernel milliseconds, launches) since the last call; CUDA events on the ctx stream."""
        ms, n = C.c_double(), C.c_int64()
        self.check(self.lib.hist_ctx_fill_time_ms(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def upload(self, profile):
        return DeviceProfile(self, profile)

    def upload_sequences(self, seqs, alphabet, components=1, wildcard="*"):
        """Leaf profiles Profile(components, alphabet, FastSeq, row) (profile.cpp:23-76) of a batch of residue
        strings (str or bytes), built on the device: only the residues cross the bus.  Tokenisation
        (fastseq.cpp:10-16) happens inside the library (hist_profile_from_residues)."""
        n = len(seqs)
        raw = [s if isinstance(s, bytes) else s.encode("latin-1", "replace") if isinstance(s, str) else bytes(s) for s in seqs]
        lens = np.fromiter(map(len, raw), np.int32, n)
        out = (C.c_void_p * n)()
        if wildcard.lower() in alphabet.lower():
            raise ValueError("the wildcard character must not be part of the alphabet")
        self.check(self.lib.hist_profile_from_residues(self.h, n, (C.c_char_p * n)(*raw), _p(lens, C.c_int32), alphabet.encode(),
                                                       components, out))
        leaf = DeviceProfile.leaf
        A = len(alphabet)
        return [leaf(self, h, s, alphabet, components, wildcard, A) for h, s in zip(out, seqs)]

    def upload_packed(self, residues, offsets, alphabet, components=1):
        """hist_profile_from_residues for a level of leaves whose residue strings sit back to back in ONE host buffer
        (residues: uint8 array; sequence k = residues[offsets[k]:offsets[k+1]]).  No per-sequence Python object is made:
        returns a HandleBatch.  What a compiled caller does with its own FASTA buffer."""
        residues = np.ascontiguousarray(residues, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.int64)
        n = len(offsets) - 1
        ptrs = (residues.ctypes.data + offsets[:-1]).astype(np.uint64)
        lens = np.diff(offsets).astype(np.int32)
        out = (C.c_void_p * n)()
        self.check(self.lib.hist_profile_from_residues(self.h, n, ptrs.ctypes.data_as(C.POINTER(C.c_char_p)), _p(lens, C.c_int32),
                                                       alphabet.encode(), components, out))
        return HandleBatch(self, out, n, "profile", keep=(residues, offsets))

    def upload_tokens(self, tokens, alph_size, components=1):
        """Same from token strings (int8 arrays, negative = wildcard): hist_profile_from_sequences."""
        toks = [np.ascontiguousarray(t, np.int8) for t in tokens]
        n = len(toks)
        lens = np.fromiter(map(len, toks), np.int32, n)
        ptrs = (C.c_void_p * n)(*[t.ctypes.data for t in toks])
        out = (C.c_void_p * n)()
        self.check(self.lib.hist_profile_from_sequences(self.h, n, ptrs, _p(lens, C.c_int32), alph_size, components, out))
        return [DeviceProfile(self, None, handle=C.c_void_p(h), seq=None, alphabet="x" * alph_size, components=components) for h in out]

    def sumprod(self, model, tree):
        return SumProd(self, model, tree)

    def close_all(self, objs):
        """Destroy a level's pairs or profiles with one call (hist_pairs_destroy / hist_profiles_destroy)."""
        if isinstance(objs, HandleBatch):
            objs.close()
            return
        objs = [o for o in objs if o.h]
        if not objs or not self.h:
            return
        fn = self.lib.hist_pairs_destroy if isinstance(objs[0], Pair) else self.lib.hist_profiles_destroy
        fn((C.c_void_p * len(objs))(*[o.h for o in objs]), len(objs))
        for o in objs:
            o.h = None

    def prepare(self, x, y, hmm, env=None):
        return Pair(self, x, y, hmm, env if env is not None else D.Envelope(-1))

    def prepare_batch(self, xs, ys, hmms, envs=None, backward=False):
        """hist_pair_prepare_batch: the DPMatrix constructors of a whole level of sibling merges.
        hmms: one PairHmm (shared) or a list; envs: list of Envelope (None = no envelope)."""
        plan = PrepareBatch(self, xs, ys, hmms, envs, backward)
        return plan.run()

    def fill_batch_async(self, pairs, backward=False, arr=None):
        """hist_fill_batch with lp_end == NULL: enqueue the fills, no read-back, no synchronisation."""
        n = len(pairs)
        if arr is None:
            arr = pairs.arr if isinstance(pairs, HandleBatch) else (C.c_void_p * n)(*[p.h for p in pairs])
        self.check(self.lib.hist_fill_batch(self.h, arr, n, int(backward), None, None))
        return arr

    def fill_results(self, pairs, backward=False):
        """hist_fill_batch_results: the likelihoods of fills enqueued with fill_batch_async (waits for the stream)."""
        n = len(pairs)
        arr = pairs.arr if isinstance(pairs, HandleBatch) else (C.c_void_p * n)(*[p.h for p in pairs])
        lp_end = np.empty(n)
        lp_start = np.empty(n) if backward else None
        self.check(self.lib.hist_fill_batch_results(self.h, arr, n, int(backward), _p(lp_end, C.c_double), _p(lp_start, C.c_double)))
        return lp_end, lp_start

    def fill_batch(self, pairs, backward=False):
        n = len(pairs)
        arr = pairs.arr if isinstance(pairs, HandleBatch) else (C.c_void_p * n)(*[p.h for p in pairs])
        lp_end = np.empty(n)
        lp_start = np.empty(n) if backward else None
        self.check(self.lib.hist_fill_batch(self.h, arr, n, int(backward), _p(lp_end, C.c_double), _p(lp_start, C.c_double)))
        return lp_end, lp_start


def diagonal_envelope(x_tok, y_tok, alph_size, sparse=True, kmer_len=6, kmer_threshold=-1, band_size=64, cell_size=40, max_size=0, lib=None):
    """DiagonalEnvelope::initFull / initSparse (diagenv.cpp:93-190) on the host: hist_diagonal_envelope.  Returns the ascending diagonals."""
    lib = lib or load()
    x, y = np.ascontiguousarray(x_tok, np.int8), np.ascontiguousarray(y_tok, np.int8)
    gp = D.GuideParams(0, 0, int(bool(sparse)), int(kmer_len), int(kmer_threshold), int(band_size), int(cell_size), 0, int(max_size))
    out, n = np.zeros(len(x) + len(y) - 1, np.int32), np.zeros(1, np.int32)
    rc = lib.hist_diagonal_envelope(None, _p(x, C.c_int8), len(x), _p(y, C.c_int8), len(y), int(alph_size), C.byref(gp), _p(out, C.c_int32), len(out), _p(n, C.c_int32))
    if rc:
        raise HistError(rc, (lib.hist_last_error(None) or b"").decode())
    return out[:int(n[0])].copy()


def align_path_merge(alignments, row_lengths, lib=None):
    """alignPathMerge (alignpath.cpp:153-207) on the host: hist_align_path_merge.  alignments: [(rows int[], bits uint8 [len(rows)][cols])];
    row_lengths: residues of every row.  Returns bits uint8 [n_rows][cols]; rows no alignment mentions stay empty."""
    lib = lib or load()
    nrows = np.array([len(r) for r, _ in alignments], np.int32)
    ncols = np.array([np.asarray(b).shape[1] if len(r) else 0 for r, b in alignments], np.int32)
    rows = np.ascontiguousarray(np.concatenate([np.asarray(r, np.int32) for r, _ in alignments] + [np.zeros(0, np.int32)]), np.int32)
    bits = np.ascontiguousarray(np.concatenate([np.asarray(b, np.uint8).ravel() for _, b in alignments] + [np.zeros(0, np.uint8)]), np.uint8)
    off = np.zeros(len(row_lengths) + 1, np.int64)
    off[1:] = np.cumsum(row_lengths)
    rcol, nout = np.zeros(max(1, int(off[-1])), np.int32), np.zeros(1, np.int32)
    rc = lib.hist_align_path_merge(None, len(alignments), _p(nrows, C.c_int32), _p(ncols, C.c_int32), _p(rows, C.c_int32), _p(bits, C.c_uint8),
                                   len(row_lengths), _p(off, C.c_int64), _p(rcol, C.c_int32), _p(nout, C.c_int32))
    if rc:
        raise HistError(rc, (lib.hist_last_error(None) or b"").decode())
    out = np.zeros((len(row_lengths), int(nout[0])), np.uint8)
    for r in range(len(row_lengths)):
        c = rcol[off[r]:off[r + 1]]
        out[r, c[c >= 0]] = 1
    return out


class HandleBatch:
    """A level's profiles or pairs as ONE array of library handles (no Python object per item): what hist_profile_from_residues
    / hist_pair_prepare_batch hand back, passed on as it is to hist_pair_prepare_batch / hist_fill_batch / hist_*_destroy."""

    def __init__(self, ctx, arr, n, kind, keep=None):
        self.ctx, self.arr, self.n, self.kind, self.keep = ctx, arr, n, kind, keep

    def __len__(self):
        return self.n

    def handles(self):
        return np.frombuffer(self.arr, np.uint64, self.n)

    def close(self):
        if self.arr is not None and self.ctx.h:
            (self.ctx.lib.hist_pairs_destroy if self.kind == "pair" else self.ctx.lib.hist_profiles_destroy)(self.arr, self.n)
        self.arr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceProfile:
    """Device-resident Profile (profile.h:41-76), see hist_profile_upload."""

    def __init__(self, ctx, profile, handle=None, seq=None, alphabet=None, components=1, wildcard="*"):
        self.ctx, self._host = ctx, profile
        self._leaf = (seq, alphabet, components, wildcard)
        if handle is None:
            d = profile.desc()
            handle = C.c_void_p()
            ctx.check(ctx.lib.hist_profile_upload(ctx.h, C.byref(d), C.byref(handle)))
        self.h = handle
        if profile is not None:
            self.components, self.alph_size = profile.components, profile.alph_size
        else:
            self.components, self.alph_size = components, len(alphabet)

    @classmethod
    def leaf(cls, ctx, handle, seq, alphabet, components, wildcard, alph_size):
        """A leaf profile the library has just built (a level holds thousands: no ctypes objects, no checks)."""
        p = object.__new__(cls)
        p.ctx, p._host, p._leaf, p.h, p.components, p.alph_size = ctx, None, (seq, alphabet, components, wildcard), handle, components, alph_size
        return p

    @property
    def host(self):
        """Host-side flattened Profile (built lazily for profiles constructed on the device)."""
        if self._host is None:
            seq, alphabet, components, wildcard = self._leaf
            seq = seq if isinstance(seq, str) else bytes(seq).decode()
            self._host = D.Profile.from_sequence(seq, alphabet, components, wildcard)
        return self._host

    def sum_path_absorb(self, log_cpt_weight, log_ins_prob, no_alloc=False):
        """Profile::calcSumPathAbsorbProbs (profile.cpp:112-131): (lpCumAbs of END, per-state lpCumAbs).
        no_alloc: through hist_profile_sum_path_absorb_noalloc (built-in device scratch, no cudaMalloc)."""
        n = self.host.n_states if self._host is not None else len(self._leaf[0]) + 2
        cum = np.empty(n)
        w = np.ascontiguousarray(log_cpt_weight, np.float64)
        ins = np.ascontiguousarray(log_ins_prob, np.float64)
        res = C.c_double()
        fn = self.ctx.lib.hist_profile_sum_path_absorb_noalloc if no_alloc else self.ctx.lib.hist_profile_sum_path_absorb
        self.ctx.check(fn(self.h, _p(w, C.c_double), _p(ins, C.c_double), _p(cum, C.c_double), C.byref(res)))
        return res.value, cum

    def close(self):
        if self.h and self.ctx.h:
            self.ctx.lib.hist_profile_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PrepareBatch:
    """Argument arrays of hist_pair_prepare_batch, built once (they only point at host buffers) and
    runnable many times: run() performs the host->device transfer and the table kernels."""

    def __init__(self, ctx, xs, ys, hmms, envs=None, backward=False):
        n = len(xs)
        self.ctx, self.xs, self.ys, self.n = ctx, list(xs), list(ys), n
        hmms = list(hmms) if isinstance(hmms, (list, tuple)) else [hmms] * n
        envs = list(envs) if envs is not None else [None] * n
        self.keep = (hmms, envs)
        uniq = {}
        self.hd = [uniq.setdefault(id(h), h.desc()) for h in hmms]
        self.ed = [None if e is None else e.desc() for e in envs]
        self.hp = (C.POINTER(D.PairHmmDesc) * n)(*[C.pointer(d) for d in self.hd])
        self.ep = (C.POINTER(D.EnvDesc) * n)(*[C.POINTER(D.EnvDesc)() if d is None else C.pointer(d) for d in self.ed])
        self.flags = 1 if backward else 0

    def run(self, xs=None, ys=None):
        """xs / ys: replacement device profiles (same order), e.g. freshly constructed leaves."""
        n, ctx = self.n, self.ctx
        if xs is not None:
            self.xs, self.ys = list(xs), list(ys)
        xp = (C.c_void_p * n)(*[x.h for x in self.xs])
        yp = (C.c_void_p * n)(*[y.h for y in self.ys])
        out = (C.c_void_p * n)()
        ctx.check(ctx.lib.hist_pair_prepare_batch(ctx.h, n, xp, yp, self.hp, self.ep, self.flags, out))
        made = Pair.made
        return [made(ctx, x, y, h) for x, y, h in zip(self.xs, self.ys, out)]

    def run_batch(self, leaves):
        """The same call for a HandleBatch of 2 n profiles, merge k joining profiles 2k and 2k + 1; returns a HandleBatch of pairs.
        The profiles must outlive the pairs (the library also holds a reference)."""
        n, ctx = self.n, self.ctx
        h = leaves.handles()
        xh, yh = np.ascontiguousarray(h[0::2]), np.ascontiguousarray(h[1::2])
        out = (C.c_void_p * n)()
        ctx.check(ctx.lib.hist_pair_prepare_batch(ctx.h, n, xh.ctypes.data_as(C.POINTER(C.c_void_p)), yh.ctypes.data_as(C.POINTER(C.c_void_p)),
                                                  self.hp, self.ep, self.flags, out))
        return HandleBatch(ctx, out, n, "pair", keep=leaves)


class Pair:
    """DPMatrix + ForwardMatrix + BackwardMatrix of one sibling merge (forward.h)."""

    def __init__(self, ctx, x, y, hmm=None, env=None, handle=None):
        self.ctx, self.x, self.y = ctx, x, y
        if handle is None:
            hd, ed = hmm.desc(), env.desc()
            handle = C.c_void_p()
            ctx.check(ctx.lib.hist_pair_prepare(ctx.h, x.h, y.h, C.byref(hd), C.byref(ed), C.byref(handle)))
        self.h = handle
        self._info = None
        self.components, self.alph_size = x.components, x.alph_size

    @classmethod
    def made(cls, ctx, x, y, handle):
        """A pair the library has just prepared as part of a batch."""
        p = object.__new__(cls)
        p.ctx, p.x, p.y, p.h, p._info, p.components, p.alph_size = ctx, x, y, handle, None, x.components, x.alph_size
        return p

    @property
    def info(self):
        """hist_pair_info, fetched on first use (a level of merges creates hundreds of pairs at once)."""
        if self._info is None:
            info = D.PairInfo()
            self.ctx.check(self.ctx.lib.hist_pair_get_info(self.h, C.byref(info)))
            self._info = info
        return self._info

    def close(self):
        if self.h and self.ctx.h:
            self.ctx.lib.hist_pair_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_cells(self):
        return int(self.info.n_cells)

    def tables(self):
        nx, ny, Cc, A = self.info.x_size, self.info.y_size, self.components, self.alph_size
        r = dict(subx=np.empty((nx, Cc, A)), suby=np.empty((ny, Cc, A)), insx=np.empty(nx), insy=np.empty(ny),
                 rootsubx=np.empty(nx), rootsuby=np.empty(ny))
        self.ctx.check(self.ctx.lib.hist_pair_fetch_tables(self.h, *[_p(r[k], C.c_double) for k in
                                                                      ("subx", "suby", "insx", "insy", "rootsubx", "rootsuby")]))
        return r

    def envelope(self, full=True):
        nx, ny = self.info.x_size, self.info.y_size
        r = dict(x_key=np.zeros(nx, np.int32), y_key=np.zeros(ny, np.int32), x_near_start=np.zeros(nx, np.uint8),
                 y_near_end=np.zeros(ny, np.uint8))
        if full:
            r["in_envelope"] = np.zeros((nx - 1, ny - 1), np.uint8)
        self.ctx.check(self.ctx.lib.hist_pair_fetch_envelope(self.h, _p(r["x_key"], C.c_int32), _p(r["y_key"], C.c_int32),
                                                             _p(r["x_near_start"], C.c_uint8), _p(r["y_near_end"], C.c_uint8),
                                                             _p(r.get("in_envelope"), C.c_uint8)))
        r["n_cells"] = self.n_cells
        return r

    def forward(self):
        lp = C.c_double()
        self.ctx.check(self.ctx.lib.hist_fwd_fill(self.h, C.byref(lp)))
        return lp.value

    def backward(self):
        lp = C.c_double()
        self.ctx.check(self.ctx.lib.hist_bwd_fill(self.h, C.byref(lp)))
        return lp.value

    def cells(self, which=0):
        """(lp [n,5], xpos [n], ypos [n]) in the order of DPMatrix::write."""
        n = self.n_cells
        lp = np.empty((max(n, 1), 5))
        xp, yp = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
        got = C.c_int64()
        self.ctx.check(self.ctx.lib.hist_cells_fetch(self.h, which, _p(lp, C.c_double), _p(xp, C.c_int32), _p(yp, C.c_int32),
                                                     n, C.byref(got)))
        return lp[:n], xp[:n], yp[:n]

    def cells_dense(self, which=0):
        """Dense [(nx-1),(ny-1),5] array with -inf where no cell exists (the oracle's layout)."""
        lp, xp, yp = self.cells(which)
        out = np.full((self.info.x_size - 1, self.info.y_size - 1, 5), -np.inf)
        out[xp, yp] = lp
        return out

    # ---- traceback (forward.cpp:257-302, 1287-1300)
    def _cells_np(self, buf, n):
        return np.frombuffer(buf, dtype=np.int32, count=3 * n).reshape(n, 3).copy()

    def trace_best(self, end=None):
        """ForwardMatrix::bestTrace(end); default end = EEE.  Returns [n,3] (xpos, ypos, state), start -> end."""
        if end is None:
            end = (self.info.x_size - 1, self.info.y_size - 1, D.EEE)
        cap = self.info.x_size + self.info.y_size + 8
        buf = (D.Cell * cap)()
        n = C.c_int64()
        self.ctx.check(self.ctx.lib.hist_trace_best(self.h, D.Cell(*[int(v) for v in end]), buf, cap, C.byref(n)))
        return self._cells_np(buf, n.value)

    def trace_best_forward(self, start):
        """BackwardMatrix::bestTrace(start): greedy walk through the Backward matrix to EEE."""
        cap = self.info.x_size + self.info.y_size + 8
        buf = (D.Cell * cap)()
        n = C.c_int64()
        self.ctx.check(self.ctx.lib.hist_trace_best_forward(self.h, D.Cell(*[int(v) for v in start]), buf, cap, C.byref(n)))
        return self._cells_np(buf, n.value)

    def trace_sample(self, mt_state, n_traces=1):
        """ForwardMatrix::sampleTrace x n_traces.  mt_state: uint32[625] (std::mt19937 image), updated in place."""
        cap = (self.info.x_size + self.info.y_size + 8) * n_traces
        buf = (D.Cell * cap)()
        off = np.zeros(n_traces + 1, np.int64)
        assert mt_state.dtype == np.uint32 and mt_state.size == 625 and mt_state.flags.c_contiguous
        self.ctx.check(self.ctx.lib.hist_trace_sample(self.h, _p(mt_state, C.c_uint32), n_traces, buf, cap, _p(off, C.c_int64)))
        allc = self._cells_np(buf, int(off[-1]))
        return [allc[off[k]:off[k + 1]] for k in range(n_traces)]

    def posterior_cells(self, min_post_prob):
        """BackwardMatrix::cellsAbovePostProbThreshold (forward.cpp:1302-1319): ([n,3] cells, [n] log posteriors)
        in priority-queue pop order."""
        n = C.c_int64()
        rc = self.ctx.lib.hist_posterior_cells(self.h, float(min_post_prob), None, None, 0, C.byref(n))
        if rc not in (0, -5):
            self.ctx.check(rc)
        cap = max(1, n.value)
        buf = (D.Cell * cap)()
        lpp = np.empty(cap)
        self.ctx.check(self.ctx.lib.hist_posterior_cells(self.h, float(min_post_prob), buf, _p(lpp, C.c_double), cap, C.byref(n)))
        return self._cells_np(buf, n.value), lpp[:n.value].copy()

    def fb_counts(self, waits):
        """Numeric part of BackwardMatrix::getCounts (forward.cpp:1183-1214): (indel[6], w_x[2,xn], w_y[2,yn],
        match cells [n,3], match weights [n]); see hist_fb_counts in include/historian_b200.h."""
        waits = np.ascontiguousarray(waits, np.float64)
        assert waits.shape == (6,)
        xn, yn = int(self.info.x_size), int(self.info.y_size)
        indel, wx, wy = np.zeros(6), np.zeros((2, xn)), np.zeros((2, yn))
        cap = max(1, int(self.info.n_cells))
        buf = (D.Cell * cap)()
        cw = np.empty(cap)
        n = C.c_int64()
        self.ctx.check(self.ctx.lib.hist_fb_counts(self.h, _p(waits, C.c_double), _p(indel, C.c_double), _p(wx, C.c_double),
                                                   _p(wy, C.c_double), buf, _p(cw, C.c_double), cap, C.byref(n)))
        return indel, wx, wy, self._cells_np(buf, n.value), cw[:n.value].copy()

    def eliminate(self, cells, strategy=D.COLLAPSE_CHAINS):
        """Numeric core of ForwardMatrix::makeProfile (forward.cpp:686-843) on a sorted cell set [n,3].
        Returns a dict: state_cell [S,3], state_lp_absorb [S,C,A], state_fwd_lp [S], trans_src/dest/lp/best_lp [T],
        via_off [T+1], via [V,3]."""
        cells = np.ascontiguousarray(cells, np.int32).reshape(-1, 3)
        n = len(cells)
        Cc, A = self.components, self.alph_size
        cap_s, cap_t, cap_v = n + 4, 8 * n + 64, 16 * n + 64
        while True:
            r = D.ElimResult()
            st_cell = (D.Cell * cap_s)()
            st_abs, st_fwd = np.empty((cap_s, Cc, A)), np.empty(cap_s)
            t_src, t_dst = np.empty(cap_t, np.int32), np.empty(cap_t, np.int32)
            t_lp, t_best = np.empty(cap_t), np.empty(cap_t)
            via_off = np.empty(cap_t + 1, np.int64)
            via = (D.Cell * cap_v)()
            r.state_cell, r.state_lp_absorb, r.state_fwd_lp = st_cell, _p(st_abs, C.c_double), _p(st_fwd, C.c_double)
            r.trans_src, r.trans_dest = _p(t_src, C.c_int32), _p(t_dst, C.c_int32)
            r.trans_lp, r.trans_best_lp = _p(t_lp, C.c_double), _p(t_best, C.c_double)
            r.trans_best_via_off, r.trans_best_via = _p(via_off, C.c_int64), via
            r.state_capacity, r.trans_capacity, r.via_capacity = cap_s, cap_t, cap_v
            rc = self.ctx.lib.hist_profile_eliminate(self.h, cells.ctypes.data_as(C.POINTER(D.Cell)), n, int(strategy), C.byref(r))
            if rc == -5 and (r.n_states > cap_s or r.n_trans > cap_t or r.n_via > cap_v):
                cap_s, cap_t, cap_v = max(cap_s, r.n_states), max(cap_t, r.n_trans), max(cap_v, int(r.n_via))
                continue
            self.ctx.check(rc)
            break
        ns, nt, nv = r.n_states, r.n_trans, int(r.n_via)
        return dict(state_cell=self._cells_np(st_cell, ns), state_lp_absorb=st_abs[:ns].copy(), state_fwd_lp=st_fwd[:ns].copy(),
                    trans_src=t_src[:nt].copy(), trans_dest=t_dst[:nt].copy(), trans_lp=t_lp[:nt].copy(),
                    trans_best_lp=t_best[:nt].copy(), via_off=via_off[:nt + 1].copy(), via=self._cells_np(via, nv))


SP_COUNTS, SP_MAP_STATES, SP_NODE_POST = 1, 2, 4


def fetch_profile(ctx, handle, with_absorb=True):
    """hist_profile_fetch: a device profile read back as a descs.Profile."""
    lib = ctx.lib
    n, nt, A, Cc = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    ctx.check(lib.hist_profile_get_sizes(handle, C.byref(n), C.byref(nt), C.byref(A), C.byref(Cc)))
    is_emit = np.empty(n.value, np.uint8)
    src, dst, lp = np.empty(nt.value, np.int32), np.empty(nt.value, np.int32), np.empty(nt.value)
    leaf = np.empty(n.value, np.int32)
    ab = np.full((n.value, Cc.value, A.value), -np.inf)
    ctx.check(lib.hist_profile_fetch(handle, _p(is_emit, C.c_uint8), _p(src, C.c_int32), _p(dst, C.c_int32), _p(lp, C.c_double),
                                     _p(ab, C.c_double) if with_absorb else None, _p(leaf, C.c_int32)))
    return D.Profile(is_emit, src, dst, lp, ab, leaf)


class Recon:
    """hist_recon: the level-batched, device-resident progressive reconstruction (recon.cpp:917-1052)."""

    def __init__(self, ctx, n_nodes, alphabet, components=1):
        self.ctx, self.n_nodes = ctx, n_nodes
        h = C.c_void_p()
        ctx.check(ctx.lib.hist_recon_create(ctx.h, n_nodes, alphabet.encode(), components, C.byref(h)))
        self.h = h

    def close(self):
        if self.h and self.ctx.h:
            self.ctx.lib.hist_recon_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_guide(self, row_nodes, bits):
        """bits: [n_rows][n_cols] bool/uint8 (True = residue); row_nodes[r] = tree node of row r."""
        bits = np.ascontiguousarray(bits, np.uint8)
        rn = np.ascontiguousarray(row_nodes, np.int32)
        self.ctx.check(self.ctx.lib.hist_recon_set_guide(self.h, bits.shape[0], bits.shape[1], _p(rn, C.c_int32), _p(bits, C.c_uint8)))

    def add_leaves(self, nodes, seqs):
        n = len(nodes)
        raw = [s if isinstance(s, bytes) else s.encode("latin-1", "replace") for s in seqs]
        lens = np.fromiter(map(len, raw), np.int32, n)
        nd = np.ascontiguousarray(nodes, np.int32)
        self.ctx.check(self.ctx.lib.hist_recon_add_leaves(self.h, n, _p(nd, C.c_int32), (C.c_char_p * n)(*raw), _p(lens, C.c_int32)))

    def level(self, merges, hmm_descs, band, strategy, profile_samples, max_profile_states, root_node=-1, mt_state=None):
        """merges: list of (node, left, right, closest_left, closest_right, closest); hmm_descs: one PairHmmDesc per merge
        (kept alive by the caller).  Returns lp_end per merge."""
        n = len(merges)
        # hist_merge_desc[n] filled column-wise (a level holds hundreds of merges): six int32 and the PairHMM pointer
        rec = np.zeros(n, np.dtype([("i", np.int32, 6), ("hmm", np.uint64)], align=True))
        assert rec.dtype.itemsize == C.sizeof(D.MergeDesc)
        rec["i"] = np.asarray(merges, np.int32).reshape(n, 6)
        addr = {}
        rec["hmm"] = [addr.setdefault(id(hd), C.addressof(hd)) for hd in hmm_descs]
        arr = rec.ctypes.data_as(C.POINTER(D.MergeDesc))
        prm = D.ReconParams()
        prm.band, prm.strategy, prm.profile_samples, prm.max_profile_states = int(band), int(strategy), int(profile_samples), int(max_profile_states)
        prm.min_post_prob = 0.0
        prm.root_node = int(root_node)
        if mt_state is not None:
            prm.mt_state = _p(mt_state, C.c_uint32)
        lp = np.empty(n)
        self.ctx.check(self.ctx.lib.hist_recon_level(self.h, n, arr, C.byref(prm), _p(lp, C.c_double)))
        return lp

    def alignment(self):
        """Dense bool [n_nodes][n_cols] of the root's best alignment path."""
        nc = C.c_int32()
        rc = self.ctx.lib.hist_recon_alignment(self.h, C.byref(nc), None, 0)
        if rc not in (0, -5):
            self.ctx.check(rc)
        bits = np.empty((self.n_nodes, max(nc.value, 0)), np.uint8)
        if bits.size:
            self.ctx.check(self.ctx.lib.hist_recon_alignment(self.h, C.byref(nc), _p(bits, C.c_uint8), bits.size))
        return bits.view(np.bool_)   # the library writes 0 / 1

    # ---- exchange between ranks (one hist_recon per GPU)
    def export_sizes(self, node):
        """(bytes of the host blob, bytes of the device table) of a live node."""
        nb, db = C.c_int64(), C.c_int64()
        rc = self.ctx.lib.hist_recon_export_node(self.h, node, None, 0, C.byref(nb), None, C.byref(db))
        if rc not in (0, -5):
            self.ctx.check(rc)
        return nb.value, db.value

    def export_node(self, node, dev_out_ptr):
        """Host blob (uint8 array) of `node`; its lpAbsorb table is copied into the device buffer at dev_out_ptr."""
        nb, db = self.export_sizes(node)
        blob = np.empty(nb, np.uint8)
        n2, d2 = C.c_int64(), C.c_int64()
        self.ctx.check(self.ctx.lib.hist_recon_export_node(self.h, node, blob.ctypes.data, nb, C.byref(n2), C.c_void_p(dev_out_ptr), C.byref(d2)))
        return blob

    def import_node(self, node, blob, dev_table_ptr):
        blob = np.ascontiguousarray(blob, np.uint8)
        self.ctx.check(self.ctx.lib.hist_recon_import_node(self.h, node, blob.ctypes.data, blob.size, C.c_void_p(dev_table_ptr)))

    def release_node(self, node):
        self.ctx.check(self.ctx.lib.hist_recon_release_node(self.h, node))

    def export_stores(self):
        nb = C.c_int64()
        rc = self.ctx.lib.hist_recon_export_stores(self.h, None, 0, C.byref(nb))
        if rc not in (0, -5):
            self.ctx.check(rc)
        blob = np.empty(nb.value, np.uint8)
        self.ctx.check(self.ctx.lib.hist_recon_export_stores(self.h, blob.ctypes.data, blob.size, C.byref(nb)))
        return blob

    def import_stores(self, blob):
        blob = np.ascontiguousarray(blob, np.uint8)
        self.ctx.check(self.ctx.lib.hist_recon_import_stores(self.h, blob.ctypes.data, blob.size))

    def stage_ms(self, reset=False):
        ms = np.zeros(len(D.RECON_STAGES))
        self.ctx.check(self.ctx.lib.hist_recon_stage_ms(self.h, _p(ms, C.c_double), int(reset)))
        return dict(zip(D.RECON_STAGES, ms.tolist()))

    def profile(self, node, with_absorb=True):
        h = self.ctx.lib.hist_recon_profile(self.h, node)
        if not h:
            raise KeyError("node %d has no live device profile" % node)
        return fetch_profile(self.ctx, C.c_void_p(h), with_absorb)


def _addr(a):
    """Raw address of a numpy array, a torch tensor (host or CUDA) or None."""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(a.ctypes.data)


class SumProd:
    """SumProduct / AlignColSumProduct (sumprod.h:32-104) for one model and tree, on the device.

    `columns` is a uint8 array [n_cols, N] (rows in the tree's postorder node order): residue
    characters, '-' or '.' for gaps, '*' (or anything outside the alphabet) for wildcards."""

    def __init__(self, ctx, model, tree):
        self.ctx = ctx
        self.C, self.A, self.N = model.components, model.alph_size, tree.n_nodes
        md, td = model.desc(), tree.desc()
        h = C.c_void_p()
        ctx.check(ctx.lib.hist_sumprod_create(ctx.h, C.byref(md), C.byref(td), C.byref(h)))
        self.h = h
        self.counts_size = int(ctx.lib.hist_sumprod_counts_size(h))

    def close(self):
        if self.h and self.ctx.h:
            self.ctx.lib.hist_sumprod_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def columns(self, columns, weights=None, flags=SP_COUNTS, counts=None, col_loglike=True):
        """The column loop of EigenCounts::accumulateSubstitutionCounts (model.cpp:900-915) over all columns.
        `counts`: flat EigenCounts buffer accumulated into (numpy array, or a CUDA torch tensor that then stays
        on the device for an all-reduce); allocated if None.  Returns a dict."""
        columns = np.ascontiguousarray(columns, np.uint8)
        L, N = columns.shape
        assert N == self.N, "columns must be [n_cols, %d]" % self.N
        r = {}
        if flags & SP_COUNTS:
            r["counts"] = np.zeros(self.counts_size) if counts is None else counts
        if col_loglike:
            r["col_loglike"] = np.zeros(L)
        if flags & SP_MAP_STATES:
            r["map_state"] = np.zeros((L, N), np.uint8)
        if flags & SP_NODE_POST:
            r["node_post"] = np.zeros((L, N, self.A))
        w = None if weights is None else np.ascontiguousarray(weights, np.float64)
        self.ctx.check(self.ctx.lib.hist_sumprod_columns(self.h, _addr(columns), _addr(w), L, int(flags), _addr(r.get("col_loglike")),
                                                         _addr(r.get("counts")), _addr(r.get("map_state")), _addr(r.get("node_post"))))
        return r

    def excluded_posteriors(self, columns, queries, normalize=True):
        """SumProduct::logNodeExcludedPostProb (sumprod.cpp:219-250) for every column and every query
        (node, excluded neighbour or -1): hist_sumprod_excluded_posteriors.  Returns [n_queries][n_cols][C][A]."""
        columns = np.ascontiguousarray(columns, np.uint8)
        L, N = columns.shape
        assert N == self.N, "columns must be [n_cols, %d]" % self.N
        nodes = np.array([q[0] for q in queries], np.int32)
        excl = np.array([q[1] for q in queries], np.int32)
        out = np.zeros((len(queries), L, self.C, self.A))
        if L:
            self.ctx.check(self.ctx.lib.hist_sumprod_excluded_posteriors(self.h, columns.ctypes.data_as(C.c_char_p), L, len(queries),
                                                                         _p(nodes, C.c_int32), _p(excl, C.c_int32), int(bool(normalize)), _p(out, C.c_double)))
        return out

    def columns_device(self, d_columns, n_cols, d_counts, d_col_loglike=None, d_weights=None, flags=SP_COUNTS):
        """Same call with every buffer already resident on the device (torch CUDA tensors)."""
        self.ctx.check(self.ctx.lib.hist_sumprod_columns(self.h, _addr(d_columns), _addr(d_weights), int(n_cols), int(flags),
                                                         _addr(d_col_loglike), _addr(d_counts), None, None))

    def debug_column(self, column):
        column = np.ascontiguousarray(column, np.uint8)
        Cc, A, N = self.C, self.A, self.N
        r = dict(E=np.zeros((Cc, N, A)), F=np.zeros((Cc, N, A)), G=np.zeros((Cc, N, A)), logE=np.zeros((Cc, N)), logF=np.zeros((Cc, N)),
                 logG=np.zeros((Cc, N)))
        ll = C.c_double()
        self.ctx.check(self.ctx.lib.hist_sumprod_debug_column(self.h, column.ctypes.data_as(C.c_char_p),
                                                              *[_p(r[k], C.c_double) for k in ("E", "F", "G", "logE", "logF", "logG")],
                                                              C.byref(ll)))
        r["col_loglike"] = ll.value
        return r

    def branch_tables(self):
        Cc, A, N = self.C, self.A, self.N
        sub, esc = np.zeros((Cc, N, A, A)), np.zeros((Cc, N, A, A, 2))
        self.ctx.check(self.ctx.lib.hist_sumprod_branch_tables(self.h, _p(sub, C.c_double), _p(esc, C.c_double)))
        return sub, esc[..., 0] + 1j * esc[..., 1]

    def transform_counts(self, counts):
        """EigenModel::getSubCounts (model.cpp:1352-1377): (rootCount[C][A], subCount[C][A][A])."""
        root, sub = np.zeros((self.C, self.A)), np.zeros((self.C, self.A, self.A))
        if not hasattr(counts, "data_ptr"):
            counts = np.ascontiguousarray(counts, np.float64)
        self.ctx.check(self.ctx.lib.hist_sumprod_transform_counts(self.h, _addr(counts), _p(root, C.c_double), _p(sub, C.c_double)))
        return root, sub

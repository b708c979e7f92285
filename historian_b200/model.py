"""Host-side substitution/indel model: the parts of the reference's RateModel / ProbModel /
EigenModel (model.h, model.cpp) that feed the device path.  Plain numpy/scipy host code:
model preparation is O(C*A^3) once per branch and is not on the accelerated path.

Mirrors (same names and argument meaning):
  RateModel.read / readComponent      model.cpp:172-236   (JSON model format of model/*.json)
  RateModel.getEqmProbVector          model.cpp:282-320   (least-squares equilibrium)
  RateModel.getSubProbMatrix          model.cpp:322-334   (matrix exponential per component)
  makeDiscretizedGammaModel           gamma.cpp:5-32
  ProbModel                           model.cpp:374-391
  IndelCounts.decayWaitTime           model.cpp:1106-1108
  EigenModel                          model.cpp:1135-1199 (eigendecomposition, inverse)
"""
import json
import math
import os

import numpy as np

from . import descs as D

MODEL_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data", "models")


class RateModel:
    def __init__(self, alphabet, wildcard="*"):
        self.alphabet = alphabet
        self.wildcard = wildcard
        self.insRate = self.delRate = 0.0
        self.insExtProb = self.delExtProb = 0.0
        self.cptWeight = []
        self.insProb = []   # [C] arrays [A]
        self.subRate = []   # [C] arrays [A][A], rows sum to zero

    @property
    def components(self):
        return len(self.subRate)

    @property
    def alphabetSize(self):
        return len(self.alphabet)

    @staticmethod
    def preset(name):
        """Named model from data/models/<name>.json (the reference embeds the same JSON as
        C strings, presets.cpp:13)."""
        with open(os.path.join(MODEL_DIR, name + ".json")) as f:
            return RateModel.read(json.load(f))

    @staticmethod
    def read(js):
        m = RateModel(js["alphabet"], js.get("wildcard", "*"))
        m.insRate, m.insExtProb = float(js["insrate"]), float(js["insextprob"])
        m.delRate, m.delExtProb = float(js["delrate"]), float(js["delextprob"])
        for cpt in (js["mixture"] if isinstance(js.get("mixture"), list) else [js]):
            m._read_component(cpt)
        w = np.array(m.cptWeight, float)
        m.cptWeight = list(w / w.sum())
        return m

    def _read_component(self, jm):
        A = self.alphabetSize
        sr = np.zeros((A, A))
        rm = jm["subrate"]
        for i, si in enumerate(self.alphabet):
            row = rm.get(si)
            if row is None:
                continue
            for j, sj in enumerate(self.alphabet):
                if j != i and sj in row:
                    sr[i, j] += float(row[sj])
                    sr[i, i] -= float(row[sj])
        if "rootprob" in jm:
            ip = np.array([float(jm["rootprob"].get(s, 0.0)) for s in self.alphabet])
        else:
            ip = self.getEqmProbVector(sr)
        self.cptWeight.append(float(jm["weight"]) if isinstance(jm.get("weight"), (int, float)) else 1.0)
        self.insProb.append(ip)
        self.subRate.append(sr)

    @staticmethod
    def getEqmProbVector(sr):
        A = sr.shape[0]
        M = np.vstack([sr.T, np.ones((1, A))])
        b = np.zeros(A + 1)
        b[A] = 1
        eqm = np.linalg.lstsq(M, b, rcond=None)[0]
        eqm = np.maximum(eqm, 0.0)
        return eqm / eqm.sum()

    def getSubProbMatrix(self, t):
        from scipy.linalg import expm
        return np.array([expm(sr * t) for sr in self.subRate])

    def gamma(self, bins, shape):
        """makeDiscretizedGammaModel (gamma.cpp:5-32)."""
        from scipy.stats import gamma as G
        assert self.components == 1
        mult = np.array([G.ppf((c + 1) / (bins + 1), a=shape, scale=shape) for c in range(bins)])
        mult /= mult.mean()
        m = RateModel(self.alphabet, self.wildcard)
        m.insRate, m.delRate, m.insExtProb, m.delExtProb = self.insRate, self.delRate, self.insExtProb, self.delExtProb
        m.cptWeight = [1.0 / bins] * bins
        m.insProb = [self.insProb[0].copy() for _ in range(bins)]
        m.subRate = [self.subRate[0] * r for r in mult]
        return m


def decayWaitTime(decayRate, timeInterval):
    return 1.0 / decayRate - timeInterval / (math.exp(decayRate * timeInterval) - 1.0)


class ProbModel:
    """ProbModel(model, t) (model.cpp:374-391)."""

    def __init__(self, model, t):
        self.t = t
        self.ins = 1 - math.exp(-model.insRate * t)
        self.del_ = 1 - math.exp(-model.delRate * t)
        self.insExt, self.delExt = model.insExtProb, model.delExtProb
        self.insWait = decayWaitTime(model.insRate, t)
        self.delWait = decayWaitTime(model.delRate, t)
        self.cptWeight = np.array(model.cptWeight)
        self.insVec = np.array(model.insProb)
        self.subMat = model.getSubProbMatrix(t)

    def as_dict(self):
        return {"ins": self.ins, "del": self.del_, "insExt": self.insExt, "delExt": self.delExt, "sub": self.subMat,
                "insProb": self.insVec, "t": self.t, "insWait": self.insWait, "delWait": self.delWait}


def pair_hmm(model, tl, tr):
    """PairHMM(ProbModel(model,tl), ProbModel(model,tr), model.insProb) as in recon.cpp:946-948."""
    l, r = ProbModel(model, tl), ProbModel(model, tr)
    return D.PairHmm.from_probs(l.as_dict(), r.as_dict(), np.array(model.insProb), np.array(model.cptWeight))


class EigenModel:
    """EigenModel(model) (model.cpp:1135-1199): R_c = V diag(lambda) V^-1 per component."""

    def __init__(self, model):
        self.model = model
        self.eval, self.evec, self.evecInv = [], [], []
        for sr in model.subRate:
            w, v = np.linalg.eig(sr)
            self.eval.append(w.astype(complex))
            self.evec.append(v.astype(complex))
            self.evecInv.append(np.linalg.inv(v.astype(complex)))

    def getSubProbMatrix(self, t):
        out = []
        for w, v, vi in zip(self.eval, self.evec, self.evecInv):
            p = (v * np.exp(w * t)[None, :]) @ vi
            out.append(np.clip(p.real, 0.0, 1.0))   # model.cpp:1272,1282
        return np.array(out)

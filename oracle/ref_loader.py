"""Import the UNMODIFIED reference modules from /root/reference (this container only).

TEST INFRASTRUCTURE ONLY — used by oracle/make_golden.py and by the oracle-vs-reference
pin tests (skipped when /root/reference is absent, as on the GPU box).

The reference imports rdkit at module import time (ggpm/nnutils.py:4-5,
ggpm/mol_graph.py:2, ggpm/chemutils.py:1-6) but the encoder arithmetic never touches it,
so rdkit is replaced by MagicMock modules.  ggpm/__init__.py star-imports everything, so a
synthetic package object with the right __path__ is registered instead of executing it.
Nothing is copied: the modules are executed from where they lie.
"""
import importlib
import os
import sys
import types
from unittest.mock import MagicMock

REFERENCE_ROOT = os.environ.get("GGPM_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "ggpm", "encoder.py"))


_cache = {}


def load_reference():
    """Returns a namespace with the reference modules: .encoder .rnn .nnutils .mol_graph .vocab"""
    if "ns" in _cache:
        return _cache["ns"]
    if not reference_available():
        raise RuntimeError("reference checkout not found at %s" % REFERENCE_ROOT)
    for n in ("rdkit", "rdkit.Chem", "rdkit.Chem.rdchem", "rdkit.Chem.AllChem",
              "rdkit.DataStructs", "rdkit.RDLogger"):
        if n not in sys.modules:
            sys.modules[n] = MagicMock()
    sys.modules["rdkit"].Chem = sys.modules["rdkit.Chem"]
    if "ggpm" not in sys.modules:
        pkg = types.ModuleType("ggpm")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "ggpm")]
        sys.modules["ggpm"] = pkg
    ns = types.SimpleNamespace()
    ns.nnutils = importlib.import_module("ggpm.nnutils")
    ns.rnn = importlib.import_module("ggpm.rnn")
    ns.mol_graph = importlib.import_module("ggpm.mol_graph")
    ns.vocab = importlib.import_module("ggpm.vocab")
    ns.encoder = importlib.import_module("ggpm.encoder")
    _cache["ns"] = ns
    return ns


class FakePairVocab:
    """Stands in for PairVocab(trio_polymer_vocab.txt): only .size() feeds the encoder ctor
    (ggpm/encoder.py:52,57); __getitem__ maps a synthetic (motif_id, attach_id) label to itself."""

    def __init__(self, n_motif=721, n_attach=6215):
        self._size = (n_motif, n_attach)

    def size(self):
        return self._size

    def __getitem__(self, x):
        return (int(x[0]), int(x[1]))


class IdAtomVocab:
    """Atom vocabulary whose labels are already integer ids (synthetic molecules)."""

    def __init__(self, n=38):
        self._n = n

    def size(self):
        return self._n

    def __getitem__(self, x):
        return int(x)

"""Import the UNMODIFIED reference modules: from /root/reference in the build container, else from the staged
byte-identical copy under oracle/_ref/ (oracle/make_ref.py; git-ignored, travels to the GPU box).

TEST INFRASTRUCTURE ONLY — used by oracle/make_golden*.py, by the oracle-vs-reference pin tests, by the
wrapper-level drop-in tests and by bench.py's reference arm / cpu_baseline leg (the one place outside tests/ that
may execute the reference).  The product (ggpm_b200/) never imports this.

The reference imports rdkit at module import time (ggpm/nnutils.py:4-5,
ggpm/mol_graph.py:2, ggpm/chemutils.py:1-6) but the encoder arithmetic never touches it,
so rdkit is replaced by MagicMock modules.  ggpm/__init__.py star-imports everything, so a
synthetic package object with the right __path__ is registered instead of executing it.
The modules are executed unmodified from where they lie (the staged copy is byte-identical, see its MANIFEST.json).
"""
import importlib
import os
import sys
import types
from unittest.mock import MagicMock

_HERE = os.path.dirname(os.path.abspath(__file__))
STAGED_ROOT = os.path.join(_HERE, "_ref")


def _resolve_root():
    for cand in (os.environ.get("GGPM_REFERENCE_ROOT"), "/root/reference", STAGED_ROOT):
        if cand and os.path.isfile(os.path.join(cand, "ggpm", "encoder.py")):
            return cand
    return os.environ.get("GGPM_REFERENCE_ROOT", "/root/reference")


REFERENCE_ROOT = _resolve_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "ggpm", "encoder.py"))


def reference_is_staged_copy() -> bool:
    return os.path.abspath(REFERENCE_ROOT) == os.path.abspath(STAGED_ROOT)


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


def load_reference_module(name: str):
    """any further module of the reference package by name ('property_vae', 'hpgnn', 'decoder',
    'property_optimizer', ...), imported unmodified under the same stub"""
    load_reference()
    return importlib.import_module("ggpm." + name)


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

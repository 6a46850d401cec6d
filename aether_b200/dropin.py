"""Make the reference's scripts pick up this implementation without editing them.

    import aether_b200.dropin; aether_b200.dropin.install()
    # ... then run experiments/lorentz/main.py from the reference tree (runpy or import)

``install()`` registers a module under the reference's dotted path ``nn.state2state.aether`` that
exports this package's ``Aether`` (experiments/lorentz/main.py:16 imports ``Aether, ParallelAether``
from there).  ``ParallelAether`` is broken as shipped in the reference (feature-count mismatch,
SURVEY.md 8a) and is exported as a class that raises on construction.
"""
import sys
import types

from .aether import Aether, AetherLocalizer, FieldNetwork
from .locs import LoCS, Localizer


class ParallelAether:  # pragma: no cover - placeholder for the reference's broken ablation
    def __init__(self, *a, **k):
        raise NotImplementedError("ParallelAether is not implemented by aether_b200 (broken in the reference)")


def install() -> types.ModuleType:
    m = types.ModuleType("nn.state2state.aether")
    m.Aether = Aether
    m.ParallelAether = ParallelAether
    m.AetherLocalizer = AetherLocalizer
    m.FieldNetwork = FieldNetwork
    m.__doc__ = "aether_b200 drop-in for the reference's nn/state2state/aether.py"
    sys.modules["nn.state2state.aether"] = m
    # the field-free baseline (experiments import it as nn.state2state.locs.locs.LoCS)
    ml = types.ModuleType("nn.state2state.locs.locs")
    ml.LoCS = LoCS
    ml.Localizer = Localizer
    ml.__doc__ = "aether_b200 drop-in for the reference's nn/state2state/locs/locs.py (LoCS only)"
    sys.modules["nn.state2state.locs.locs"] = ml
    # the FiLM / attention-pooled field variant (experiments/lorentz/main.py: --model film_aether)
    from .dynamic_field import DynamicFieldAether, LatentFieldNetwork
    md = types.ModuleType("nn.state2state.dynamic_field_aether")
    md.DynamicFieldAether = DynamicFieldAether
    md.LatentFieldNetwork = LatentFieldNetwork
    md.__doc__ = "aether_b200 drop-in for the reference's nn/state2state/dynamic_field_aether.py"
    sys.modules["nn.state2state.dynamic_field_aether"] = md
    return m

"""rotors_mpc_controller_b200 -- B200-native drop-in for the acados-backed NMPC of
Anaskherro/rotors_mpc_controller (same exports as the reference's ``__init__.py:3-12``)."""
from .closed_loop import ClosedLoop, Plant
from .controller import BatchedPositionNMPC, ControllerParams, PositionNMPC
from .params import (DYNAMIC_RANGES, apply_dynamic_configuration,
                     clamp_dynamic_configuration, load_params)
from .reference import BatchReference, ReferenceGenerator, stack_reference

__all__ = [
    'PositionNMPC',
    'BatchedPositionNMPC',
    'ClosedLoop',
    'Plant',
    'ControllerParams',
    'ReferenceGenerator',
    'BatchReference',
    'stack_reference',
    'apply_dynamic_configuration',
    'clamp_dynamic_configuration',
    'DYNAMIC_RANGES',
    'load_params',
]

from enum import IntEnum, auto


class Phase(IntEnum):
    """FedOBD state machine (/root/reference/simulation_lib/method/fed_obd/phase.py:4-7)."""

    STAGE_ONE = auto()
    STAGE_TWO = auto()
    END = auto()

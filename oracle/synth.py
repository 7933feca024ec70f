"""Re-export of peft_b200.synth (deterministic synthetic weights / inputs) for the oracle and the tests.
TEST INFRASTRUCTURE: the factories themselves contain no model arithmetic and live in the package so that bench.py's own
arm does not import anything from oracle/."""
from peft_b200.synth import *  # noqa: F401,F403
from peft_b200.synth import VitCfg, make_batch, make_head, make_state_dict, trainable_names  # noqa: F401

from .cartpole import CartpoleEnvCfg
from .coverage import OpcodeCoverageEnvCfg
from .velocity import AnymalCFlatEnvCfg, AnymalCRoughEnvCfg, H1RoughEnvCfg, UnitreeGo1RoughEnvCfg

TASKS = {
    "Isaac-Cartpole-v0": CartpoleEnvCfg,
    "Isaac-Velocity-Flat-Anymal-C-v0": AnymalCFlatEnvCfg,
    "Isaac-Velocity-Rough-Anymal-C-v0": AnymalCRoughEnvCfg,
    "Isaac-Velocity-Rough-Unitree-Go1-v0": UnitreeGo1RoughEnvCfg,
    "Isaac-Velocity-Rough-H1-v0": H1RoughEnvCfg,
    "B200-Opcode-Coverage-v0": OpcodeCoverageEnvCfg,   # synthetic: opcodes / actuator kinds the five tasks above do not use
}

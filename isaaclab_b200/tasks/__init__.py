from .cartpole import CartpoleEnvCfg
from .coverage import OpcodeCoverageEnvCfg
from .velocity import (AnymalBFlatEnvCfg, AnymalBRoughEnvCfg, AnymalCFlatEnvCfg, AnymalCRoughEnvCfg, AnymalDFlatEnvCfg,
                       AnymalDRoughEnvCfg, CassieFlatEnvCfg, CassieRoughEnvCfg, G1FlatEnvCfg, G1RoughEnvCfg, H1FlatEnvCfg,
                       H1RoughEnvCfg, UnitreeA1FlatEnvCfg, UnitreeA1RoughEnvCfg, UnitreeGo1FlatEnvCfg,
                       UnitreeGo1RoughEnvCfg, UnitreeGo2FlatEnvCfg, UnitreeGo2RoughEnvCfg, play)

TASKS = {
    "Isaac-Cartpole-v0": CartpoleEnvCfg,
    "Isaac-Velocity-Flat-Anymal-C-v0": AnymalCFlatEnvCfg,
    "Isaac-Velocity-Rough-Anymal-C-v0": AnymalCRoughEnvCfg,
    "Isaac-Velocity-Rough-Unitree-Go1-v0": UnitreeGo1RoughEnvCfg,
    "Isaac-Velocity-Rough-H1-v0": H1RoughEnvCfg,
    # the other quadrupeds of TASKS/.../locomotion/velocity/config: same managers, other robot constants
    "Isaac-Velocity-Flat-Anymal-B-v0": AnymalBFlatEnvCfg,
    "Isaac-Velocity-Rough-Anymal-B-v0": AnymalBRoughEnvCfg,
    "Isaac-Velocity-Flat-Anymal-D-v0": AnymalDFlatEnvCfg,
    "Isaac-Velocity-Rough-Anymal-D-v0": AnymalDRoughEnvCfg,
    "Isaac-Velocity-Flat-Unitree-A1-v0": UnitreeA1FlatEnvCfg,
    "Isaac-Velocity-Rough-Unitree-A1-v0": UnitreeA1RoughEnvCfg,
    "Isaac-Velocity-Flat-Unitree-Go1-v0": UnitreeGo1FlatEnvCfg,
    "Isaac-Velocity-Flat-Unitree-Go2-v0": UnitreeGo2FlatEnvCfg,
    "Isaac-Velocity-Rough-Unitree-Go2-v0": UnitreeGo2RoughEnvCfg,
    # the other bipeds
    "Isaac-Velocity-Flat-H1-v0": H1FlatEnvCfg,
    "Isaac-Velocity-Flat-G1-v0": G1FlatEnvCfg,
    "Isaac-Velocity-Rough-G1-v0": G1RoughEnvCfg,
    "Isaac-Velocity-Flat-Cassie-v0": CassieFlatEnvCfg,
    "Isaac-Velocity-Rough-Cassie-v0": CassieRoughEnvCfg,
    "B200-Opcode-Coverage-v0": OpcodeCoverageEnvCfg,   # synthetic: opcodes / actuator kinds the tasks above do not use
}
# "<id minus -v0>-Play-v0": the 50-env evaluation variant of every task above (velocity.play)
for _id, _fn in list(TASKS.items()):
    if _id.startswith("Isaac-Velocity-"):
        TASKS[_id[:-3] + "-Play-v0"] = (lambda fn=_fn: play(fn()))

"""garfield_b200 — Byzantine-resilient distributed training for B200 (sm_100a).

Sub-packages: `aggregators` (GAR registry), `garfieldpp` (Server / Worker / Byz* roles, tools, datasets), `models`,
`ops` (drivers of the hand-written kernels: `gar`, `optim`, `attacks`, `nn`), `parallel` (flat and peer-mapped arenas,
worker fusion, the SSMW / MSMW / decentralized device engines), `native` (C++ host rules), `compat` (TensorFlow-side
counterparts), `utils` (the `tools` toolbox, checkpoints, tracing), `apps` (the trainers behind `applications/`).
Nothing heavy is imported here: the native libraries are built and loaded on first use (`garfield_b200._build`).
"""

__version__ = "0.1.0"

__all__ = ["aggregators", "apps", "compat", "garfieldpp", "models", "native", "ops", "parallel", "utils", "__version__"]

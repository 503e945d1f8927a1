"""obst: B200-native implementation of the OBST (HomebrewNLP-MTF) mixer training-step hot path.

Host side mirrors the reference's public surface for that path -- ModelParameter / BlockArgs (src/dataclass.py),
LAYER_FUNCTIONS (src/model/frontend.py), build (src/model/__init__.py), OPTIMIZERS / get_optimizer
(src/optimizer), MODULES (src/optimizer/learning_rate.py), get_train_model (src/run/train.py) -- but executes
eagerly on one B200 per process through the C-ABI kernels of include/obst_b200.h.
"""
__version__ = "0.1.0"

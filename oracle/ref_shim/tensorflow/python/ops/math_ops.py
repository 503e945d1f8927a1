import numpy as np

sign = lambda x, name=None: np.sign(x)  # noqa: E731

import numpy as np

reshape = lambda x, shape, name=None: np.reshape(x, shape)  # noqa: E731
diag_part = lambda x, name=None: np.diagonal(x).copy()  # noqa: E731
matrix_transpose = lambda x, name=None: np.swapaxes(x, -1, -2)  # noqa: E731

"""The scalars a training step logs, under the reference's names, and a TensorBoard event file for them.

The reference exports `learning_rate, first_loss, last_loss, mean_loss` (+ `token_loss` with language inputs,
`accuracy` with calc_accuracy) from `get_train_model`'s return values and writes each as a TF2 `summary.scalar` at the
global step (src/run/run.py:123-132,145; src/run/utils_run.py:32-43).  `log_dict` builds the same dictionary from
this package's `get_train_model`; `SummaryWriter` writes it as `events.out.tfevents.*` in the format TensorBoard reads
-- TFRecord-framed `Event` protos whose values are rank-0 float tensors tagged for the `scalars` plugin, as
`tf.summary.scalar` emits them -- without TensorFlow: the few protobuf fields are encoded by hand (event.proto,
summary.proto, tensor.proto).  Host-side only; one device->host read per distinct loss object."""
import os
import socket
import struct
import time
import typing

from ..dataclass import ModelParameter
from ..inputs import masked_crc32c
from ..model import Loss

DT_FLOAT = 1
DATA_CLASS_SCALAR = 1


def log_dict(params: ModelParameter, out: typing.Sequence) -> typing.Dict[str, float]:
    """src/run/run.py:123-132 on the tuple `get_train_model` returned."""
    _, _, learning_rate, loss, first_loss, last_loss, video_loss, token_loss, accuracy, _, _ = out
    cache: typing.Dict[int, typing.Tuple[float, float, float]] = {}

    def values(obj: Loss):
        if id(obj) not in cache:
            cache[id(obj)] = Loss.values(obj.sums.cpu(), obj.n_global, obj.z_loss)
        return cache[id(obj)]

    lr = float(learning_rate.item() if hasattr(learning_rate, "item") else learning_rate)
    scalars = {"learning_rate": lr, "first_loss": values(first_loss)[0], "last_loss": values(last_loss)[0],
               "mean_loss": values(loss)[0]}
    if params.use_video and video_loss is not None:
        scalars["video_loss"] = values(video_loss)[0]
    if params.use_language:
        scalars["token_loss"] = values(token_loss)[0]
    if accuracy is not None:
        scalars["accuracy"] = values(accuracy)[2]
    return scalars


# ---------------------------------------------------------------------------------------------------------
# protobuf wire format, the handful of fields an event file needs
# ---------------------------------------------------------------------------------------------------------
def _varint(x: int) -> bytes:
    x &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _len(field: int, payload: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def scalar_value(tag: str, value: float, with_metadata: bool = True) -> bytes:
    """Summary.Value{tag = 1, metadata = 9, tensor = 8}: a rank-0 DT_FLOAT TensorProto, plugin `scalars`."""
    tensor = _varint((1 << 3) | 0) + _varint(DT_FLOAT) + _len(2, b"") + _len(5, struct.pack("<f", value))
    out = _len(1, tag.encode("utf-8")) + _len(8, tensor)
    if with_metadata:
        metadata = _len(1, _len(1, b"scalars")) + _varint((4 << 3) | 0) + _varint(DATA_CLASS_SCALAR)
        out += _len(9, metadata)
    return out


def encode_event(wall_time: float, step: int, *, file_version: typing.Optional[str] = None,
                 values: typing.Sequence[bytes] = ()) -> bytes:
    """Event{wall_time = 1 (double), step = 2 (int64), file_version = 3 | summary = 5 {value = 1 ...}}."""
    out = _varint((1 << 3) | 1) + struct.pack("<d", wall_time)
    if step:
        out += _varint((2 << 3) | 0) + _varint(step)
    if file_version is not None:
        out += _len(3, file_version.encode())
    if values:
        out += _len(5, b"".join(_len(1, v) for v in values))
    return out


class SummaryWriter:
    """`summary.create_file_writer(logdir)` + `summary.scalar(...)` + `summary.flush()` for host floats."""

    def __init__(self, logdir: str, filename_suffix: str = ".v2"):
        os.makedirs(logdir, exist_ok=True)
        now = time.time()
        self.path = os.path.join(logdir, f"events.out.tfevents.{int(now):010d}.{socket.gethostname()}.{os.getpid()}"
                                         f"{filename_suffix}")
        self._file = open(self.path, "ab")
        self._seen: typing.Set[str] = set()  # TensorFlow's writer keeps a tag's metadata on its first event only
        self._record(encode_event(now, 0, file_version="brain.Event:2"))

    def _record(self, payload: bytes):
        header = struct.pack("<Q", len(payload))
        self._file.write(header + struct.pack("<I", masked_crc32c(header)) + payload +
                         struct.pack("<I", masked_crc32c(payload)))

    def scalar(self, tag: str, value: float, step: int, wall_time: typing.Optional[float] = None):
        first = tag not in self._seen
        self._seen.add(tag)
        self._record(encode_event(time.time() if wall_time is None else wall_time, int(step),
                                  values=[scalar_value(tag, float(value), first)]))

    def scalars(self, values: typing.Dict[str, float], step: int):
        """One step's `log_dict`, each key an event of its own as the reference writes them."""
        for tag, value in values.items():
            self.scalar(tag, value, step)

    def flush(self):
        self._file.flush()

    def close(self):
        self._file.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

"""In-order event tapes (include/viekf_batch.h: viekf_event / viekf_tape).

The tape replaces the reference's rewind/replay queue (vi_ekf_meas.cpp:6-194): it is the in-order sequence of
propagate_state / update / init_feature / keep_only_features calls that the queue ends up applying.
"""
import numpy as np

from . import abi


class TapeBuilder:
    def __init__(self):
        self._ev = []
        self._ids = []
        self._n_value_blocks = 0

    def _new(self, kind):
        e = np.zeros((), dtype=abi.EVENT_DTYPE)
        e["kind"] = kind
        e["id"] = -1
        e["noise"] = -1
        self._ev.append(e)
        return e

    def value_block(self):
        """Reserve one 8-double per-instance value block; returns its index."""
        k = self._n_value_blocks
        self._n_value_blocks += 1
        return k

    @property
    def n_value_blocks(self):
        return self._n_value_blocks

    def propagate(self, u, t, per_instance=False, sigma=None, sample=-1):
        """propagate_state(u, t, save_input=False) (reference vi_ekf.cpp:262-318).  sigma: per-instance white noise
        std of the 6 IMU channels, drawn on the device when the tape is run with a seed."""
        e = self._new(abi.EV_PROP)
        e["t"] = t
        e["aux"] = sample
        if u is not None:
            e["z"][:6] = u
        if sigma is not None:
            e["sigma"][:6] = sigma
        if per_instance:
            e["noise"] = self.value_block()
        return int(e["noise"])

    def update(self, mtype, z, R, active=True, id=-1, depth=float("nan"), per_instance=False, sigma=None, from_imu=False):
        """add_measurement + update, in order (reference vi_ekf_meas.cpp:130-147,196-278)."""
        e = self._new(abi.EV_UPD)
        z = np.atleast_1d(np.asarray(z, dtype=np.float64))
        if sigma is not None:
            sg = np.atleast_1d(np.asarray(sigma, dtype=np.float64))
            e["sigma"][: sg.size] = sg
        R = np.atleast_2d(np.asarray(R, dtype=np.float64))
        e["type"] = mtype
        e["zdim"] = z.size
        e["rdim"] = R.shape[0]
        e["z"][: z.size] = z
        e["z"][4] = depth
        e["R"][: R.size] = R.T.ravel()  # column-major
        e["active"] = int(active)
        e["id"] = id
        e["aux"] = 1 if from_imu else 0
        if per_instance:
            e["noise"] = self.value_block()
        return int(e["noise"])

    def init_feature(self, pixel, id, depth=float("nan"), per_instance=False):
        e = self._new(abi.EV_INIT)
        e["z"][:2] = pixel
        e["z"][4] = depth
        e["id"] = id
        if per_instance:
            e["noise"] = self.value_block()
        return int(e["noise"])

    def clear_feature(self, id):
        e = self._new(abi.EV_CLEAR)
        e["id"] = id

    def keep_only_features(self, ids):
        e = self._new(abi.EV_KEEP)
        e["ids_off"] = len(self._ids)
        e["n_ids"] = len(ids)
        self._ids.extend(int(i) for i in ids)

    def keyframe_reset(self):
        self._new(abi.EV_KFRESET)

    def stat(self, bucket, x_true17, per_instance=False):
        """Accumulate error statistics against the true base state (17 entries, reference state layout)."""
        e = self._new(abi.EV_STAT)
        x_true17 = np.asarray(x_true17, dtype=np.float64)
        e["id"] = bucket
        e["z"][:8] = x_true17[:8]
        e["R"][:9] = x_true17[8:17]
        if per_instance:
            e["noise"] = self.value_block()
        return int(e["noise"])

    def events(self):
        if not self._ev:
            return np.zeros(0, dtype=abi.EVENT_DTYPE)
        return np.array(self._ev, dtype=abi.EVENT_DTYPE)

    def id_pool(self):
        return np.asarray(self._ids if self._ids else [0], dtype=np.int32)

    def n_imu_steps(self):
        return sum(1 for e in self._ev if int(e["kind"]) == abi.EV_PROP)

"""
Asynchronous trajectory writer.

The reference snapshots every colloid through Python property reads at each write interval, BEFORE the chunk is
integrated (/root/reference/swarmrl/engine/espresso.py:1110-1130, 1278-1285), keeps `write_chunk_size` frames in
`traj_holder` and appends them to `trajectory.hdf5` (datasets `Times (T,1,1)`, `Ids / Types (T,N,1)`,
`Unwrapped_Positions / Velocities / Directors (T,N,3)`; espresso.py:1054-1108, 1132-1159).  Here the hot loop never
waits for any of that:

    run stream    : swarm_snapshot_dev -> one kernel writes the frame into a slot of a DEVICE ring, event
    copy stream   : waits for that event, cudaMemcpyAsync device slot -> PINNED host slot, event
    writer thread : waits for the copy event, books the frame into the pending chunk -- the pending arrays are VIEWS of
                    the pinned slots, a frame is not copied again on the host -- and writes a chunk file whenever
                    `write_chunk_size` frames are pending or every slot is taken (HDF5 in the reference's layout when
                    h5py is importable, numbered `.npz` chunks with the same arrays otherwise); writing frees the slots

The ring holds `write_chunk_size` frames when that fits the pinned-memory budget (1 GiB), else as many as fit (chunks
are then shorter).  `snapshot()` blocks only when every slot is taken and its chunk is still being written -- at that
point the rollout produces frames faster than the file system takes them (C4 with a frame per slice is 10 GB/s).
`flush()` drains the ring and writes the partial chunk without taking a snapshot (espresso.py:1310-1318, `finalize`).
"""
from __future__ import annotations

import ctypes as C
import queue
import threading

import numpy as np

from . import _lib

KEYS = ("Times", "Ids", "Types", "Unwrapped_Positions", "Velocities", "Directors")


class AsyncTrajectoryWriter:
    def __init__(self, swarm, types, out_folder, group_tag, chunk_size, ring=None, h5_filename=None):
        torch = swarm.torch
        self.swarm, self.torch = swarm, torch
        self.n = int(swarm.n)
        self.types = np.asarray(types, dtype=np.int64).copy()
        self.ids = np.arange(self.n, dtype=np.int64)
        self.out_folder, self.group_tag, self.chunk_size = out_folder, group_tag, int(chunk_size)
        self.h5_filename = h5_filename
        frame_bytes = 72 * max(self.n, 1)
        if ring is None:
            ring = max(2, min(self.chunk_size, (1 << 30) // frame_bytes))
        self.ring = int(ring)
        m = max(self.n, 1)
        self.dev_ring = min(self.ring, 4)           # a device slot is free again as soon as its frame has been copied out
        self.dev = torch.empty((self.dev_ring, 3, m, 3), dtype=torch.float64, device=swarm.dev)  # [slot][pos|vel|dir][n][3]
        self.dev_copied = [None] * self.dev_ring
        self.pin = torch.empty((self.ring, 3, m, 3), dtype=torch.float64).pin_memory()
        self.copy_stream = torch.cuda.Stream(device=swarm.dev)
        self.copied = [None] * self.ring            # event: the D2H copy of the slot has finished
        self.free = threading.Semaphore(self.ring)  # slots whose frame has been written to a file
        self.q = queue.Queue()
        self.pending = {k: [] for k in KEYS}        # the reference's traj_holder
        self.frames_taken = 0
        self.frames_written = 0
        self.chunks_written = 0
        self.error = None
        try:
            import h5py
        except ModuleNotFoundError:
            h5py = None
        self._h5py = h5py
        if h5py is not None and self.h5_filename is not None:
            self._init_h5()
        self.thread = threading.Thread(target=self._work, name="swarm-traj-writer", daemon=True)
        self.thread.start()

    # ------------------------------------------------------------------ hot-loop side
    def snapshot(self, time):
        """Frame of the CURRENT device state at simulation time `time`; returns as soon as the work is enqueued."""
        if self.error is not None:
            raise self.error
        self.free.acquire()                          # back-pressure: only when every slot is still in flight
        slot = self.frames_taken % self.ring
        torch, sw = self.torch, self.swarm
        run = torch.cuda.current_stream(sw.dev)
        dslot = self.frames_taken % self.dev_ring
        if self.dev_copied[dslot] is not None:
            run.wait_event(self.dev_copied[dslot])   # the device slot's previous frame has left the device
        d = self.dev[dslot]
        _lib.check(sw.lib.swarm_snapshot_dev(sw.h, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), C.c_void_p(run.cuda_stream)))
        taken = torch.cuda.Event()
        taken.record(run)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(taken)
            self.pin[slot].copy_(d, non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.copy_stream)
        self.copied[slot] = done
        self.dev_copied[dslot] = done
        self.frames_taken += 1
        self.q.put((slot, float(time), done))

    # ------------------------------------------------------------------ writer thread
    def _work(self):
        while True:
            item = self.q.get()
            try:
                if item is None:
                    return
                if item[0] == "flush":
                    self._write_chunk()
                    continue
                slot, time, done = item
                done.synchronize()
                f = self.pin[slot].numpy()
                self.pending["Times"].append(np.array([[time]]))
                self.pending["Ids"].append(self.ids[:, None])
                self.pending["Types"].append(self.types[:, None])
                self.pending["Unwrapped_Positions"].append(f[0, : self.n])   # views of the pinned slot
                self.pending["Velocities"].append(f[1, : self.n])
                self.pending["Directors"].append(f[2, : self.n])
                if len(self.pending["Times"]) >= min(self.chunk_size, self.ring):
                    self._write_chunk()
            except Exception as exc:  # surfaced by the next snapshot() / flush()
                self.error = exc
                for _ in range(self.ring):
                    self.free.release()
            finally:
                self.q.task_done()

    def _init_h5(self):
        n, L = self.n, self.chunk_size
        with self._h5py.File(str(self.h5_filename), "a") as f:
            g = f.require_group(self.group_tag)
            kw = dict(compression="gzip")
            g.require_dataset("Times", shape=(L, 1, 1), maxshape=(None, 1, 1), dtype=float, **kw)
            for name in ("Ids", "Types"):
                g.require_dataset(name, shape=(L, n, 1), maxshape=(None, n, 1), dtype=int, **kw)
            for name in ("Unwrapped_Positions", "Velocities", "Directors"):
                g.require_dataset(name, shape=(L, n, 3), maxshape=(None, n, 3), dtype=float, **kw)

    def _write_chunk(self):
        n_new = len(self.pending["Times"])
        if n_new == 0:
            return
        arrays = {k: np.stack(v, axis=0) for k, v in self.pending.items()}
        if self._h5py is not None and self.h5_filename is not None:
            with self._h5py.File(str(self.h5_filename), "a") as f:
                g = f[self.group_tag]
                for key, val in arrays.items():
                    ds = g[key]
                    ds.resize(self.frames_written + n_new, axis=0)
                    ds[self.frames_written: self.frames_written + n_new, ...] = val
        else:
            fn = self.out_folder / f"trajectory.{self.group_tag}.{self.chunks_written:05d}.npz"
            # large swarms: gzip would be the bottleneck of the whole rollout (72 B per particle and frame)
            (np.savez if arrays["Unwrapped_Positions"].nbytes > (64 << 20) else np.savez_compressed)(fn, **arrays)
        self.chunks_written += 1
        self.frames_written += n_new
        for v in self.pending.values():
            v.clear()
        for _ in range(n_new):      # the slots of the written frames may be overwritten now
            self.free.release()

    # ------------------------------------------------------------------ control
    def drain(self):
        """Wait until every snapshot taken so far sits in the pending chunk (or in a file)."""
        self.q.join()
        if self.error is not None:
            raise self.error

    def flush(self):
        """finalize(): everything pending goes to disk; no snapshot is taken."""
        self.q.put(("flush",))
        self.drain()

    def close(self):
        if self.thread.is_alive():
            self.q.put(None)
            self.thread.join(timeout=30)

#!/usr/bin/env python
"""End-to-end fit from JPEG files on local disk (the reference's real input path, client_fit_model.py:19-43,54-90):
sorted *.jpg listing -> shuffle(1337) -> Generator (cv2.imread + cvtColor + resize, mask > 0) -> LocalTrainer.fit on the
engine.  Compares decode inline in the fit loop (what the reference does) with the batch-ahead prefetch thread
(cfg.data.prefetch_batches) and the per-batch worker pool (cfg.data.workers).  Prints one JSON line."""
import argparse, json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--dir", default="")
    a = ap.parse_args()
    import cv2
    from crackfl.config import fast_test_config
    from crackfl.data.generator import gen_train_val_data, synth_crack_batch
    from crackfl.models.spec import REFERENCE_SPEC as S
    from crackfl.trainer import LocalTrainer
    root = a.dir or tempfile.mkdtemp(prefix="crackfl_jpeg_")
    img_dir, mask_dir = os.path.join(root, "images"), os.path.join(root, "masks")
    os.makedirs(img_dir, exist_ok=True); os.makedirs(mask_dir, exist_ok=True)
    t0 = time.perf_counter()
    for i in range(0, a.n, 32):
        x, y = synth_crack_batch(min(32, a.n - i), a.size, a.size, 4242 + i)
        for j in range(x.shape[0]):
            cv2.imwrite(os.path.join(img_dir, f"{i + j:06d}.jpg"), cv2.cvtColor(x[j], cv2.COLOR_RGB2BGR))
            cv2.imwrite(os.path.join(mask_dir, f"{i + j:06d}.jpg"), y[j, ..., 0] * 255)
    write_s = time.perf_counter() - t0
    def cgroup_cpus():
        try:
            q, per = open("/sys/fs/cgroup/cpu.max").read().split()
            return None if q == "max" else float(q) / float(per)
        except Exception:
            return None

    def decode_rate(nthreads):
        """Pure decode throughput (no GPU): every batch of the dataset once, whole batches on `nthreads` threads."""
        from concurrent.futures import ThreadPoolExecutor
        ds, _ = gen_train_val_data(img_dir, mask_dir, a.batch, (a.size, a.size), count=a.n, workers=0)
        t = time.perf_counter()
        if nthreads <= 1:
            for i in range(len(ds)):
                ds[i]
        else:
            with ThreadPoolExecutor(nthreads) as ex:
                list(ex.map(ds.__getitem__, range(len(ds))))
        return len(ds) * a.batch / (time.perf_counter() - t)

    out = {"files": a.n, "size": a.size, "batch": a.batch, "epochs": a.epochs, "cpus": os.cpu_count(), "write_seconds": write_s,
           "sched_affinity_cpus": len(os.sched_getaffinity(0)), "cgroup_cpu_quota": cgroup_cpus(),
           "decode_only_images_per_sec": {str(k): decode_rate(k) for k in (1, 4, 8)}, "arms": []}
    for label, workers, prefetch, threads in (("inline decode (reference behaviour)", 0, 0, 0),
                                              ("per-image worker pool x4, no prefetch", 4, 0, 0),
                                              ("prefetch 8 batches, 1 thread", 0, 8, 1), ("prefetch 8 batches, 4 threads", 0, 8, 4),
                                              ("prefetch 16 batches, 8 threads", 0, 16, 8),
                                              ("prefetch 32 batches, 14 threads", 0, 32, 14)):
        cfg = fast_test_config()
        cfg.train.backend = "b200"; cfg.train.batch_size = a.batch; cfg.train.image_size = (a.size, a.size)
        cfg.train.local_epochs = a.epochs; cfg.train.tensorboard = False; cfg.train.validate = False
        cfg.client.log_root = ""; cfg.data.prefetch_batches = prefetch; cfg.data.prefetch_threads = threads
        tr_ds, va_ds = gen_train_val_data(img_dir, mask_dir, a.batch, (a.size, a.size), count=a.n, workers=workers)
        tr = LocalTrainer(cfg, train_ds=tr_ds, val_ds=None, log=lambda *s: None)
        w, m = tr.fit(S.init_weights(0), 1)
        out["arms"].append({"arm": label, "workers": workers, "prefetch_batches": prefetch, "prefetch_threads": threads,
                            "images_per_sec": m["images_per_sec"], "fit_seconds": m["fit_seconds"], "loss": m["loss"]})
    print(json.dumps(out))


if __name__ == "__main__":
    main()

"""``model_evaluate`` - the server-side module the reference imports (fl_server.py:12) but does not ship.

Contract inferred from its four call sites:

    eval_model = evaluate_LocalModel(16, 224, class_names)      fl_server.py:230
    made_model = eval_model.buildGlobalModel(3, 0.001)           fl_server.py:231
    eval_model.get_weights(made_model) -> bytes                  fl_server.py:24   (client does pickle.loads on it)
    eval_model.train_model_tosave(made_model) -> {'loss','accuracy'}   fl_server.py:31-32 (dead call site)

The arguments are leftovers of a 10-class MobileNetV2 project; the global model has to be the clients' U-Net for
``set_weights`` to work (112 tensors), so that is what is built.  ``get_weights`` returns the *aggregated* weights
when ``./server_weights/weights.pickle`` exists (fl_server.py:104-105 writes it and nothing visible reads it back:
quirk Q2), which is what makes FedAvg actually feed back.
"""
import os

import numpy as np

from ..config import Config
from ..io import weights as wio
from ..models.spec import REFERENCE_SPEC


class GlobalModel(object):
    def __init__(self, spec, weights, lr):
        self.spec, self.weights, self.lr = spec, weights, lr

    def get_weights(self):
        return self.weights

    def set_weights(self, w):
        self.weights = [np.asarray(a, np.float32) for a in w]


class evaluate_LocalModel(object):
    def __init__(self, batch_size=16, img_size=224, class_names=None, cfg: Config = None, val_ds=None):
        self.batch_size, self.img_size, self.class_names = batch_size, img_size, class_names
        self.cfg, self.val_ds = cfg or Config.load(os.environ.get("CRACKFL_CONFIG")), val_ds
        self.spec = REFERENCE_SPEC

    def buildGlobalModel(self, channels=3, lr=0.001, seed=0):
        return GlobalModel(self.spec, self.spec.init_weights(seed), lr)

    def get_weights(self, model, weights_file="./server_weights/weights.pickle"):
        if os.path.exists(weights_file):
            model.set_weights(wio.load_pickle_file(weights_file))
        return wio.dumps_pickle(model.get_weights())

    def train_model_tosave(self, model):
        """Server-side evaluation of the global model (the reference's ``trainNextRound`` hook, fl_server.py:27-37)."""
        from ..trainer import make_backend
        if self.val_ds is None:
            from ..data.generator import make_datasets
            _, self.val_ds = make_datasets(self.cfg)
        be = make_backend(self.cfg, self.spec)
        be.begin_round(model.get_weights())
        be.reset_metrics()
        for i in range(len(self.val_ds)):
            be.eval_batch(*self.val_ds[i])
        m = be.read_metrics()
        return {"loss": m["loss"], "accuracy": m["accuracy"]}

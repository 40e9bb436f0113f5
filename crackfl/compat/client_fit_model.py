"""``client_fit_model`` - ``learning_fit(mt, ec, bs, pr, cr)`` with the reference's call surface
(client_fit_model.py:46-52,225-240; used at fl_client.py:102,109,154,165)."""
import os

from ..config import Config
from ..data.generator import Generator  # noqa: F401  (the reference module also exports it)
from ..io import weights as wio
from ..trainer import LocalTrainer

DEFAULT_CONFIG = None  # set to a crackfl.config.Config to change data dirs / backend for reference-style callers


class learning_fit(object):
    def __init__(self, mt, ec, bs, pr, cr):
        self.model_type, self.epochs, self.batch_size, self.round = mt, ec, bs, cr
        self.params = wio.loads_pickle(pr) if isinstance(pr, (bytes, bytearray)) else list(pr)
        cfg = DEFAULT_CONFIG or Config.load(os.environ.get("CRACKFL_CONFIG"))  # also honours CRACKFL_<SECTION>__<KEY>
        cfg.transport.allow_pickle = True
        self.cfg = cfg
        self.trainer = LocalTrainer(cfg)

    def manage_train(self, params=None, cr=None):
        """Fits from ``params`` (pickled bytes, as the server sends them) or from the constructor's parameters, writes
        ``./saved_weight/weights.pickle`` like the reference (client_fit_model.py:239-240) and - unlike it - also
        returns the weights."""
        if self.params == list():
            return []
        start = wio.loads_pickle(params) if params is not None else self.params
        w, _ = self.trainer.fit(start, cr if cr is not None else self.round)
        wio.save_pickle_file(self.cfg.client.weights_path, w)
        return w

    def Predict(self, model=None):
        from ..postproc.measure import predict_and_measure
        return predict_and_measure(self.trainer.backend.predict_batch, self.trainer.predict_ds or self.trainer.val_ds,
                                   self.cfg.measure)

"""Surrogate cost model for model-based tuning: gradient-boosted regression trees on the flattened numeric config.
The reference uses xgboost's rank:pairwise objective (tuner/cost_model.py); scikit-learn's GradientBoostingRegressor on
max-normalised targets plays that role here (xgboost is optional and used when importable)."""
import numpy as np


class XGBoostCostModel(object):

    def __init__(self, loss_type, num_threads=None, log_interval=25, upper_model=None):
        assert loss_type in ("reg", "rank")
        self.loss_type = loss_type
        self.model = None

    def fit(self, xs, ys):
        x_train = np.array(xs, dtype=np.float32)
        y_train = np.array(ys, dtype=np.float32)
        y_train = y_train / max(float(np.max(y_train)), 1e-9)
        if len(x_train) < 2:
            self.model = ("const", float(y_train.mean()) if len(y_train) else 0.0)
            return
        try:
            import xgboost as xgb
            params = {"max_depth": 3, "gamma": 0.0001, "min_child_weight": 1, "subsample": 1.0, "eta": 0.3, "lambda": 1.0, "alpha": 0,
                      "objective": "reg:squarederror" if self.loss_type == "reg" else "rank:pairwise", "verbosity": 0}
            self.model = ("xgb", xgb.train(params, xgb.DMatrix(x_train, y_train), num_boost_round=50))
        except ImportError:
            from sklearn.ensemble import GradientBoostingRegressor
            m = GradientBoostingRegressor(n_estimators=50, max_depth=3, learning_rate=0.3, subsample=1.0, random_state=0)
            m.fit(x_train, y_train)
            self.model = ("sk", m)

    def predict(self, xs):
        x = np.array(xs, dtype=np.float32)
        kind, m = self.model
        if kind == "const":
            return np.full(len(x), m, dtype=np.float32)
        if kind == "xgb":
            import xgboost as xgb
            return m.predict(xgb.DMatrix(x))
        return m.predict(x)

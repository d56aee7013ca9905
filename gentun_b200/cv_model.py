"""Model plugins: the interface the reference calls GentunModel (gentun/models/generic_models.py:7-18)
and its xgboost implementation (gentun/models/xgboost_models.py:11-37), re-hosted on the B200 engine."""
from . import engine as _engine


class GentunModel(object):
    """A model that can score itself on (x_train, y_train) by n-fold cross-validation."""

    def __init__(self, x_train, y_train):
        self.x_train = x_train
        self.y_train = y_train

    def cross_validate(self):
        raise NotImplementedError("Use a subclass with a defined model.")


class XgboostModel(GentunModel):
    """Same constructor and return value as the reference's XgboostModel; the n-fold boosted-tree
    training that xgb.cv did on host cores runs as CUDA kernels instead."""

    def __init__(self, x_train, y_train, hyperparameters, booster='gbtree', objective='reg:linear',
                 eval_metric='rmse', nfold=5, num_boost_round=5000, early_stopping_rounds=100):
        super(XgboostModel, self).__init__(x_train, y_train)
        # the reference dict.update()s the genes into its parameter dict (xgboost_models.py:16-22)
        self.params = {'booster': booster, 'objective': objective, 'eval_metric': eval_metric, 'silent': 1}
        self.params.update(hyperparameters)
        self.eval_metric = eval_metric
        self.nfold = nfold
        self.num_boost_round = num_boost_round
        self.early_stopping_rounds = early_stopping_rounds

    def fit_parameters(self):
        return {
            'booster': self.params['booster'], 'objective': self.params['objective'],
            'eval_metric': self.eval_metric, 'nfold': self.nfold,
            'num_boost_round': self.num_boost_round, 'early_stopping_rounds': self.early_stopping_rounds,
        }

    def cross_validate(self):
        """Mean validation metric over the folds (last row of xgb.cv's table, xgboost_models.py:37)."""
        return _engine.evaluate_many(self.x_train, self.y_train, [self.params], self.fit_parameters())[0]

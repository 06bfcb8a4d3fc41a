from .robust_aggregation import RobustAggregator, is_weight_param, vectorize_weight

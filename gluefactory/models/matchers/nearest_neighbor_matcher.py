"""`get_model('matchers.nearest_neighbor_matcher')` -> the B200 nearest-neighbour matcher
(see glue_factory_b200.models.nearest_neighbor_matcher)."""
from glue_factory_b200.models.nearest_neighbor_matcher import NearestNeighborMatcher as _NearestNeighborMatcher

__main_model__ = _NearestNeighborMatcher

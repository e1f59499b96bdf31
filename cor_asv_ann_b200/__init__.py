"""B200-native neural hot path of cor-asv-ann behind the reference's Sequence2Sequence API."""
GAP = '\a'

"""B200-native lockstep Hnefatafl engine."""

"""B200-native drop-in for the WorldMirror aggregator hot path."""

"""B200-native (sm_100a) implementation of the GDev96/SuperResolution HybridSuperResolutionModel hot path."""

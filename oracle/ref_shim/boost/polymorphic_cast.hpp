// named by orb_feature.cpp:4, unused

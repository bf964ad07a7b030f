// boost::span is named by feature_interface.hpp:7 but not used on the hot path

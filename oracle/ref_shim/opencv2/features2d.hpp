#include "../mslam_cv_shim.hpp"

#include "../../../mslam_eigen_shim.hpp"

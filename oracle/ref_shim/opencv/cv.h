// TEST INFRASTRUCTURE: forwards to the cv shim (oracle/ref_shim/opencv2/core/core.hpp).
#pragma once
#include "opencv2/core/core.hpp"

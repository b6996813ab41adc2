// TEST INFRASTRUCTURE — see opencv2/core/core.hpp in this shim
#pragma once
#include "../core/core.hpp"

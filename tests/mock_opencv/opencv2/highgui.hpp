// mock: everything lives in opencv2/core.hpp (see its header comment)
#pragma once
#include <opencv2/core.hpp>

// mock of the legacy umbrella header include/ORBextractor.h:26 pulls in
#pragma once
#include <opencv2/core.hpp>

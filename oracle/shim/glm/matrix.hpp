// forwarding header of the oracle glm shim (see glm.hpp)
#pragma once
#include <glm/glm.hpp>

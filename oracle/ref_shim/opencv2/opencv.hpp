// TEST INFRASTRUCTURE (oracle/_ref): forwarding header, see core/core.hpp
#pragma once
#include "core/core.hpp"

// GSL shim (oracle infrastructure, not product code): see gslshim.h
#pragma once
#include "../gslshim.h"

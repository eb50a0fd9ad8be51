// Forwarding header of the expat-backed Xerces-C compatibility shim (test infrastructure; see ../shim.hpp).
#pragma once
#include <xercesc/shim.hpp>

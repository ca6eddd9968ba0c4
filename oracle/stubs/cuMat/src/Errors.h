// TEST INFRASTRUCTURE ONLY. Empty stand-in for the one cuMat header that the reference's volume.cpp includes
// (renderer/volume.cpp:6) and uses nothing of; the cuMat submodule is absent from /root/reference.
#pragma once

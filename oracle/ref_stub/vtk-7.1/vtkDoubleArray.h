// TEST INFRASTRUCTURE -- see vtk_recording.h
#pragma once
#include "vtk_recording.h"

/* Forwarding header: the reference includes <opencv2/features2d/features2d.hpp>; everything it uses is in cvstub.hpp (test infrastructure). */
#include "../../cvstub.hpp"

/* Forwarding header: the reference includes <opencv2/imgproc/imgproc.hpp>; everything it uses is in cvstub.hpp (test infrastructure). */
#include "../../cvstub.hpp"

// RayTracer.h - host-side mirror of the reference's RayTracer class (reference
// ray/src/RayTracer.h:26-75) over the C-ABI of include/rt_b200.h.  Same member names, argument
// meaning and error behaviour; the bodies forward to the device path.  The settings the reference
// reads from its global `traceUI` (ui/TraceUI.h:35-129) are the public `ui` member (an rt_params).
//
// glm is not a dependency of this tree: dvec3 is three doubles with the same layout as glm::dvec3.
#pragma once
#include <string>
#include <vector>
#include "rt_b200.h"

namespace rtb200 {

struct dvec3 {
	double x, y, z;
};
struct ray {
	dvec3 p, d;
};

class RayTracer {
public:
	explicit RayTracer(int device = 0);
	// Renders on devices 0 .. n-1 of the box (rt_multi_*: scene replicated, interleaved tiles); call before loadScene.
	// The reference's counterpart is its thread count (-t, main.cpp:12): parallelism inside traceImage.
	bool useDevices(int n);
	~RayTracer();
	RayTracer(const RayTracer&) = delete;
	RayTracer& operator=(const RayTracer&) = delete;

	// reference RayTracer.cpp:196-240.  Returns false after reporting the message through alert().
	bool loadScene(const char* fn);
	bool sceneLoaded() const { return loaded; }
	// TraceUI::smartLoadCubemap's result: six face images, +x -x +y -y +z -z; call before loadScene
	void setCubemap(const std::vector<std::string>& faces) { cubemap = faces; }

	void traceSetup(int w, int h);             // RayTracer.cpp:242-267
	void traceImage(int w, int h);             // :279-314 (synchronous, like checkRender() == true)
	dvec3 tracePixel(int i, int j);            // :81-102
	dvec3 traceRay(ray& r, double thresh, int depth, double& length, dvec3* curr_m_kt = nullptr); // :108-174
	int aaImage() { return 0; }                // :367-370
	bool checkRender() { return true; }        // :372-375
	void waitRender() {}                       // :377-381

	dvec3 getPixel(int i, int j);              // :383-387
	void setPixel(int i, int j, dvec3 color);  // :388-394
	void getBuffer(unsigned char*& buf, int& w, int& h); // :184-189
	double aspectRatio();                      // :191-194

	const rt_stats& lastStats() const { return stats; }
	const std::string& lastAlert() const { return alert_msg; }

	rt_params ui;    // the traceUI settings that cross the boundary
	bool stopTrace;

private:
	void alert(const std::string& msg);
	rt_ctx* ctx;
	rt_multi* multi; // non-null after useDevices(n > 1): ctx is then its first device's context
	void* arena;
	size_t arena_bytes;
	bool loaded;
	std::vector<std::string> cubemap;
	std::vector<unsigned char> buffer;
	int buffer_width, buffer_height;
	rt_stats stats;
	std::string alert_msg;
};

} // namespace rtb200

// oracle/shim: headless stand-in for the reference's pbrt::GUI
// (/root/reference/src/pbrt/util/gui.h:27; gui.cpp needs glfw/glad/OpenGL, and
// there is no graphics stack on either box - SURVEY.md F14).  Only reached
// under --interactive, which the oracle never passes.  TEST INFRASTRUCTURE ONLY.
#include <pbrt/util/gui.h>
#include <cstdio>
#include <cstdlib>
namespace pbrt {
static void nogui() { fprintf(stderr, "oracle shim: GUI unavailable on this box\n"); abort(); }
GUI::GUI(std::string, Vector2i, Bounds3f) { nogui(); }
GUI::~GUI() {}
DisplayState GUI::RefreshDisplay() { nogui(); return EXIT; }
void GUI::Initialize() { nogui(); }
Point2i GUI::GetResolution() { nogui(); return Point2i(0, 0); }
}

// hw1-hw5 text scene parser.  Grammar and quirks per SURVEY appendix B; behaviour follows
// hw1/src/sceneio.cpp:8-97, hw2/src/sceneio.cpp:8-150, hw4/src/sceneio.cpp:8-108,
// hw5/src/sceneio.cpp:8-97:
//  * line oriented, first token is the keyword, trailing garbage ignored;
//  * a failed numeric extraction leaves the remaining fields untouched (operator>> semantics);
//  * inside NEW_PRIMITIVE / NEW_LIGHT the first non-matching line is re-dispatched at top level;
//  * unknown keyword => "UNKNOWN COMMAND: <kw>" on stderr, unknown shape => "UNKNWOWN FIGURE".
#include <fstream>
#include <iostream>
#include <sstream>
#include "hmath.h"
#include "host_scene.h"

namespace b200rt_host {
namespace {

// operator>> on three floats: stops at the first failure, later fields keep their values.
void read3(std::istream &in, float *v) { in >> v[0] >> v[1] >> v[2]; }

struct TextParser {
    std::istream &in;
    b200rt_host_scene &s;
    int mode;               // 0 = infer
    bool saw_samples = false, saw_light = false, saw_triangle = false;

    // hw1-hw4: the shape line must directly follow NEW_PRIMITIVE; hw5: anywhere in the block.
    bool parse_shape(const std::string &name, std::istream &ss, b200rt_primitive &p, bool &known) {
        known = true;
        if (name == "ELLIPSOID") { p.type = B200RT_PRIM_ELLIPSOID; read3(ss, p.data); }
        else if (name == "PLANE") { p.type = B200RT_PRIM_PLANE; read3(ss, p.data); }
        else if (name == "BOX") { p.type = B200RT_PRIM_BOX; read3(ss, p.data); }
        else if (name == "TRIANGLE") {
            // hw5/src/sceneio.cpp: three vertices stored as data3, data2, data
            p.type = B200RT_PRIM_TRIANGLE; saw_triangle = true;
            read3(ss, p.data3); read3(ss, p.data2); read3(ss, p.data);
        } else known = false;
        return known;
    }

    bool parse_attr(const std::string &cmd, std::istream &ss, b200rt_primitive &p) {
        if (cmd == "COLOR") read3(ss, p.color);
        else if (cmd == "POSITION") read3(ss, p.position);
        else if (cmd == "ROTATION") ss >> p.rotation[0] >> p.rotation[1] >> p.rotation[2] >> p.rotation[3];
        else if (cmd == "METALLIC") p.material = B200RT_MAT_METALLIC;
        else if (cmd == "DIELECTRIC") p.material = B200RT_MAT_DIELECTRIC;
        else if (cmd == "IOR") ss >> p.ior;
        else if (cmd == "EMISSION") read3(ss, p.emission);
        else return false;
        return true;
    }

    // Returns true and fills `next` when a line must be re-dispatched.
    bool primitive(std::string &next) {
        b200rt_primitive p{};
        p.rotation[3] = 1.f;  // identity (hw1/src/include/primitives.h:26, hw3+ quaternion.h:26)
        p.ior = 1.f;          // uninitialised in the reference when absent; unused then
        std::string line;
        if (!getline(in, line)) { return false; }
        bool known = false;
        {
            std::stringstream ss(line);
            std::string name;
            ss >> name;
            parse_shape(name, ss, p, known);
            if (!known) {
                std::stringstream ss2(line);
                std::string cmd; ss2 >> cmd;
                if (!parse_attr(cmd, ss2, p)) {
                    std::cerr << "UNKNWOWN FIGURE: " << name << '@' << line << std::endl;
                    next = line;   // hw1 drops the figure; we drop it in every mode
                    return true;
                }
            }
        }
        bool redispatch = false;
        while (getline(in, line)) {
            std::stringstream ss(line);
            std::string cmd;
            ss >> cmd;
            if (parse_attr(cmd, ss, p)) continue;
            bool k2 = false;
            if (!known && parse_shape(cmd, ss, p, k2)) { known = true; continue; }
            next = line; redispatch = true;
            break;
        }
        if (known) {
            s.prims.push_back(p);
        }
        return redispatch;
    }

    // hw2/src/sceneio.cpp:59-99: fields in any order; LIGHT_DIRECTION makes it directional.
    bool light(std::string &next) {
        b200rt_light l{};
        float direction[3] = {0, 0, 0}, position[3] = {0, 0, 0};
        bool directed = false, redispatch = false;
        std::string line;
        while (getline(in, line)) {
            std::stringstream ss(line);
            std::string cmd; ss >> cmd;
            if (cmd == "LIGHT_INTENSITY") read3(ss, l.intensity);
            else if (cmd == "LIGHT_POSITION") read3(ss, position);
            else if (cmd == "LIGHT_DIRECTION") { read3(ss, direction); directed = true; }
            else if (cmd == "LIGHT_ATTENUATION") read3(ss, l.attenuation);
            else { next = line; redispatch = true; break; }
        }
        l.type = directed ? B200RT_LIGHT_DIRECTIONAL : B200RT_LIGHT_POINT;
        const float *src = directed ? direction : position;
        l.v[0] = src[0]; l.v[1] = src[1]; l.v[2] = src[2];
        s.lights.push_back(l);
        return redispatch;
    }

    void run() {
        b200rt_scene_desc &d = s.desc;
        d.samples = 1; d.ray_depth = 1;
        std::string line;
        while (getline(in, line)) {
            while (true) {
                std::stringstream ss(line);
                std::string cmd;
                ss >> cmd;
                std::string next;
                bool again = false;
                if (cmd == "DIMENSIONS") ss >> d.width >> d.height;
                else if (cmd == "BG_COLOR") read3(ss, d.bg_color);
                else if (cmd == "CAMERA_POSITION") read3(ss, d.cam_pos);
                else if (cmd == "CAMERA_RIGHT") read3(ss, d.cam_right);
                else if (cmd == "CAMERA_UP") read3(ss, d.cam_up);
                else if (cmd == "CAMERA_FORWARD") read3(ss, d.cam_forward);
                else if (cmd == "CAMERA_FOV_X") ss >> d.cam_fov;
                else if (cmd == "RAY_DEPTH") ss >> d.ray_depth;
                else if (cmd == "SAMPLES") { ss >> d.samples; saw_samples = true; }
                else if (cmd == "AMBIENT_LIGHT") { read3(ss, d.ambient); saw_light = true; }
                else if (cmd == "NEW_LIGHT") { saw_light = true; again = light(next); }
                else if (cmd == "NEW_PRIMITIVE") again = primitive(next);
                else if (cmd != "") std::cerr << "UNKNOWN COMMAND: " << cmd << std::endl;
                if (again) { line = next; continue; }
                break;
            }
        }
    }
};

}  // namespace

int load_text_scene(const char *path, int32_t mode, b200rt_host_scene &s) {
    std::ifstream in(path);
    if (!in) return B200RT_HOST_ERR_IO;
    double t0 = now_ms();
    TextParser p{in, s, mode};
    p.run();
    if (mode == 0) {
        if (p.saw_triangle) mode = B200RT_MODE_HW5;
        else if (p.saw_samples) mode = B200RT_MODE_HW4;
        else if (p.saw_light) mode = B200RT_MODE_HW2;
        else mode = B200RT_MODE_HW1;
    }
    if (mode < B200RT_MODE_HW1 || mode > B200RT_MODE_HW5) return B200RT_ERR_INVALID;
    s.desc.mode = mode;
    if (mode == B200RT_MODE_HW1) { s.desc.samples = 1; s.desc.ray_depth = 1; }
    if (mode == B200RT_MODE_HW2) s.desc.samples = 1;
    // plane normal normalised at load in hw2/hw3/hw4 only
    if (mode == B200RT_MODE_HW2 || mode == B200RT_MODE_HW3 || mode == B200RT_MODE_HW4)
        for (auto &pr : s.prims)
            if (pr.type == B200RT_PRIM_PLANE) {
                V3 n = V3(pr.data[0], pr.data[1], pr.data[2]).normalized();
                pr.data[0] = n.x; pr.data[1] = n.y; pr.data[2] = n.z;
            }
    if (s.desc.width <= 0 || s.desc.height <= 0) return B200RT_HOST_ERR_PARSE;
    s.parse_ms = now_ms() - t0;
    if (mode == B200RT_MODE_HW5) {       // hw5/src/sceneio.cpp:92-93: initBVH, then initDistribution
        double t1 = now_ms();
        build_figure_accel(s);
        s.bvh_ms = now_ms() - t1;
    }
    s.sync();
    return B200RT_OK;
}

}  // namespace b200rt_host

"""A second, independent reading of the reference's path-tracing shader, in scalar Python.  TEST INFRASTRUCTURE ONLY.

Written from src/shaders/ComputeShader.wgsl of the reference (CS below) function by function -- oracle/crt_oracle.c was
consulted only for WHERE DESIGN.md section 3 places an fmaf, not for what the shader does.  It parses the binding buffers by
the WGSL struct layout rules itself (CS:22-47), keeps the shader's structure (one Python function per WGSL function, the same
local names, dictionaries for the structs) and runs one invocation of main() (CS:77-117) at a time.  tests/test_wgsl_literal.py compares it with the C oracle bit for bit: a mistake in the oracle's restatement of the
shader's semantics (operand order, which branch resets what, which index a light is looked up by, rand() call order ...)
would have to be made twice, in two languages and two program structures, to go unseen.  The reference itself still cannot
be executed on this box (no WGSL implementation), so the oracle stays "unpinned by the reference"; this narrows what that
leaves open to the arithmetic both readings take from DESIGN.md section 3:

* f32 everywhere: numpy float32 scalars (one IEEE rounding per + - * / sqrt);
* where section 3 says "fmaf", `Machine.fma` -- libm's fmaf through ctypes when `fused`, else two roundings (the other
  evaluation WGSL allows; the second test shows the choice moves no result by more than rounding noise unless a path
  changes its hit sequence);
* sin / cos / exp are passed in (the tests hand over include/crt_detmath.h's, whose accuracy test_oracle_kat.py bounds).

The triangle primitive (category 2) is this project's extension: not in the shader, not here.
"""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np

f32 = np.float32
_libm = C.CDLL("libm.so.6")
_libm.fmaf.restype = C.c_float
_libm.fmaf.argtypes = [C.c_float, C.c_float, C.c_float]
_libm.tan.restype = C.c_double
_libm.tan.argtypes = [C.c_double]

# CS:11-20
PI = f32(3.14159265359)
INFINITY = f32(0x7F800000)          # the integer 2139095040 converted to f32, not +inf
MAX_U32_VALUE = 0xFFFFFFFF
MAXDEPTH = 100
GRID_SIZE = 16
lambda_min = f32(400.0)
lambda_max = f32(700.0)
DIFFUSE, LIGHT, GLASS = 0, 1, 2
M32 = 0xFFFFFFFF


def u32_of(x) -> int:
    """WGSL u32(f32): truncation toward zero, out-of-range values clamp, NaN -> 0 (the pin of DESIGN section 3)."""
    x = float(x)
    if not (x > 0.0):
        return 0
    if x >= 4294967296.0:
        return M32
    return int(x)


class Machine:
    def __init__(self, sincos, exp, pow_1_2p4=None, fused: bool = True):
        self.fused, self._sincos, self._exp, self._pow = fused, sincos, exp, pow_1_2p4

    def fma(self, a, b, c):
        if self.fused:
            return f32(_libm.fmaf(float(a), float(b), float(c)))
        return f32(f32(a) * f32(b)) + f32(c)

    def sincos(self, x):
        s, c = self._sincos(f32(x))
        return f32(s), f32(c)

    def exp(self, x):
        return f32(self._exp(f32(x)))

    def pow_1_2p4(self, x):
        """pow(x, 1.0/2.4) of gamma_correct (CS:449): the exponent is the f32 constant 0.41666666."""
        return f32(self._pow(f32(x)))


# ------------------------------------------------------------------ vectors (tuples of f32)
def v_add(a, b): return tuple(x + y for x, y in zip(a, b))
def v_sub(a, b): return tuple(x - y for x, y in zip(a, b))
def v_neg(a): return tuple(-x for x in a)
def v_scale(a, s): return tuple(x * s for x in a)
def v_div(a, s): return tuple(x / s for x in a)
def v_mul(a, b): return tuple(x * y for x, y in zip(a, b))


class Shader:
    """One shader module bound to one set of buffers (the bind group of CS:1-9)."""

    def __init__(self, bindings: dict, m: Machine, max_depth: int = MAXDEPTH, grid_size: int = GRID_SIZE):
        self.m = m
        self.MAXDEPTH, self.GRID_SIZE = max_depth, grid_size
        cam = struct.unpack("<14f", bytes(bindings[4])[:56])                     # Camera CS:22-29: vec3 at 0, 16, 32; f32 at 44, 48, 52
        self.camera = dict(eye=tuple(map(f32, cam[0:3])), lookat=tuple(map(f32, cam[4:7])), up=tuple(map(f32, cam[8:11])),
                           viewport_width=f32(cam[11]), viewport_height=f32(cam[12]), focal_length=f32(cam[13]))
        cie = np.frombuffer(bytes(bindings[5]), dtype="<f4")
        self.CIE = cie.reshape(3, 471)                                           # array<array<f32,471>,3> CS:6
        sp = np.frombuffer(bytes(bindings[6]), dtype="<f4")
        self.spectra = sp.reshape(-1, 301)                                       # array<array<f32,301>> CS:7
        self.lights = self._primitives(bytes(bindings[7]))
        self.primitives = self._primitives(bytes(bindings[8]))
        self.reset_statistics()

    @staticmethod
    def _primitives(raw: bytes):
        """Primitive CS:41-47: u32 at 0; vec3<f32> (align 16) at 16, 32, 48; vec4<u32> at 64; stride 80."""
        out = []
        for o in range(0, len(raw), 80):
            cat, = struct.unpack_from("<I", raw, o)
            d1 = tuple(map(f32, struct.unpack_from("<3f", raw, o + 16)))
            d2 = tuple(map(f32, struct.unpack_from("<3f", raw, o + 32)))
            d3 = tuple(map(f32, struct.unpack_from("<3f", raw, o + 48)))
            d4 = struct.unpack_from("<4I", raw, o + 64)
            out.append(dict(category=cat, data1=d1, data2=d2, data3=d3, data4=d4))
        return out

    # -------------------------------------------------------------- built-ins with their section-3 evaluation
    def dot(self, a, b):
        m = self.m
        return m.fma(a[2], b[2], m.fma(a[1], b[1], a[0] * b[0]))

    def cross(self, a, b):
        m = self.m
        return (m.fma(a[1], b[2], -(a[2] * b[1])), m.fma(a[2], b[0], -(a[0] * b[2])), m.fma(a[0], b[1], -(a[1] * b[0])))

    def length(self, a):
        return np.sqrt(self.dot(a, a))

    def normalize(self, a):
        return v_div(a, self.length(a))

    def a_plus_s_b(self, a, s, b):
        """a + s * b per lane (ray_at CS:306, sample_light CS:353, the film point CS:493)."""
        return tuple(self.m.fma(s, y, x) for x, y in zip(a, b))

    def reflect(self, e1, e2):
        """WGSL reflect(e1, e2) = e1 - 2 * dot(e2, e1) * e2."""
        k = f32(2.0) * self.dot(e2, e1)
        return self.a_plus_s_b(e1, -k, e2)

    def refract(self, e1, e2, e3):
        """WGSL refract: k = 1 - e3^2 (1 - dot(e2,e1)^2); k < 0 -> 0 vector; else e3*e1 - (e3*dot(e2,e1) + sqrt(k)) * e2."""
        d = self.dot(e2, e1)
        k = f32(1.0) - (e3 * e3) * (f32(1.0) - d * d)
        if k < 0:
            return (f32(0), f32(0), f32(0))
        s = self.m.fma(e3, d, np.sqrt(k))
        return self.a_plus_s_b(v_scale(e1, e3), -s, e2)

    # -------------------------------------------------------------- RNG CS:865-899
    @staticmethod
    def tea(val0: int, val1: int) -> int:
        v0, v1, s0 = val0 & M32, val1 & M32, 0
        for _ in range(16):
            s0 = (s0 + 0x9e3779b9) & M32
            v0 = (v0 + ((((v1 << 4) & M32) + 0xa341316c) & M32 ^ ((v1 + s0) & M32) ^ (((v1 >> 5) + 0xc8013ea4) & M32))) & M32
            v1 = (v1 + ((((v0 << 4) & M32) + 0xad90777d) & M32 ^ ((v0 + s0) & M32) ^ (((v0 >> 5) + 0x7e95761e) & M32))) & M32
        return v0

    def pcg4d(self):
        x, y, z, w = [(s * 1664525 + 1013904223) & M32 for s in self.seed]
        x = (x + y * w) & M32
        y = (y + z * x) & M32
        z = (z + x * y) & M32
        w = (w + y * z) & M32
        x, y, z, w = x ^ (x >> 16), y ^ (y >> 16), z ^ (z >> 16), w ^ (w >> 16)
        x = (x + y * w) & M32
        y = (y + z * x) & M32
        z = (z + x * y) & M32
        w = (w + y * z) & M32
        self.seed = [x, y, z, w]

    def rand(self):
        self.pcg4d()
        self.n_rand += 1
        return f32(self.seed[0] & 0x00ffffff) / f32(0x01000000)

    # -------------------------------------------------------------- intersection CS:503-721
    @staticmethod
    def create_intersection_context(exclude: int, t_min=f32(0.001)):
        zero = (f32(0), f32(0), f32(0))
        return dict(t_min=t_min, t_max=INFINITY, index=MAX_U32_VALUE, exclude=exclude, hit=False, ray_origin=zero, ray_direction=zero)

    def ray_at(self, ray, t):
        return self.a_plus_s_b(ray["origin"], t, ray["direction"])

    def ray_intersection(self, primitive, ray, context, shape_intersection):
        category = primitive["category"]
        self.prim_tests += category in (0, 1)
        if category == 0:
            exclude = context["exclude"]
            index = primitive["data4"][3]
            if exclude == index:
                return
            edge1 = primitive["data2"]
            edge2 = primitive["data3"]
            normal = self.normalize(self.cross(edge1, edge2))
            direction = ray["direction"]
            ndotd = self.dot(normal, direction)
            if ndotd > 0:
                normal = v_neg(normal)
            ndotd = self.dot(normal, direction)
            if abs(ndotd) < f32(0.0001):
                return
            origin = ray["origin"]
            origin_to_origin = v_sub(primitive["data1"], origin)
            t = self.dot(normal, origin_to_origin) / ndotd
            t_min, t_max = context["t_min"], context["t_max"]
            if t < t_min or t > t_max:
                return
            intersection = self.ray_at(ray, t)
            m = v_sub(intersection, primitive["data1"])
            u = self.dot(m, edge1) / self.dot(edge1, edge1)
            v = self.dot(m, edge2) / self.dot(edge2, edge2)
            if u < 0.0 or u > 1.0 or v < 0.0 or v > 1.0:
                return
            shape_intersection.update(position=intersection, normal=normal, emission_index=primitive["data4"][0],
                                      reflectance_index=primitive["data4"][1], material=primitive["data4"][2])
            context.update(t_max=t, index=index, hit=True, ray_origin=origin, ray_direction=direction)
            return
        if category == 1:
            exclude = context["exclude"]
            index = primitive["data4"][3]
            if exclude == index:
                return
            center = primitive["data1"]
            radius = primitive["data2"][0]
            radius_squared = radius * radius
            origin = ray["origin"]
            direction = ray["direction"]
            center_to_origin = v_sub(origin, center)
            a = self.dot(direction, direction)
            b = f32(2.0) * self.dot(direction, center_to_origin)
            c = self.dot(center_to_origin, center_to_origin) - radius_squared
            discriminant = self.m.fma(b, b, -((f32(4.0) * a) * c))
            if discriminant <= 0.0:
                return
            t = (-b - np.sqrt(discriminant)) / (f32(2.0) * a)
            t_min, t_max = context["t_min"], context["t_max"]
            if t < t_min or t > t_max:
                t = (-b + np.sqrt(discriminant)) / (f32(2.0) * a)
                if t < t_min or t > t_max:
                    return
            intersection = self.ray_at(ray, t)
            normal = self.normalize(v_sub(intersection, center))
            shape_intersection.update(position=intersection, normal=normal, emission_index=primitive["data4"][0],
                                      reflectance_index=primitive["data4"][1], material=primitive["data4"][2])
            context.update(t_max=t, index=primitive["data4"][3], hit=True, ray_origin=origin, ray_direction=direction)

    def intersect(self, ray, exclude: int, shadow: bool):
        zero = (f32(0), f32(0), f32(0))
        shape_intersection = dict(position=zero, normal=zero, emission_index=0, reflectance_index=0, material=0)   # `var` zero value
        intersection_context = self.create_intersection_context(exclude, self.t_min)
        for primitive in self.primitives:
            if shadow and primitive["data4"][2] == GLASS:
                continue
            self.ray_intersection(primitive, ray, intersection_context, shape_intersection)
        return dict(shape_intersection=shape_intersection, context=intersection_context)

    def shadow_intersect(self, ray, include: int, exclude: int):
        self.rays_shadow += 1
        intersection = self.intersect(ray, exclude, False)
        if intersection["context"]["index"] != include:
            intersection["context"]["hit"] = False
        return intersection

    # -------------------------------------------------------------- spectra CS:309-338
    def sample_spectrum(self, index: int, lambdas):
        row = self.spectra[min(index, self.spectra.shape[0] - 1)]                 # out-of-bounds read clamps (section 3)
        return tuple(f32(row[l]) for l in lambdas)

    def sample_wavelengths(self):
        u = self.rand()
        rng = u32_of(lambda_max - lambda_min) + 1
        lam = u32_of(self.mix(f32(0), lambda_max - lambda_min + f32(1), u))
        return (lam, (lam + 4) % rng, (lam + 8) % rng, (lam + 12) % rng)

    @staticmethod
    def mix(e1, e2, e3):
        return e1 * (f32(1.0) - e3) + e2 * e3

    # -------------------------------------------------------------- lights CS:340-408
    def sample_lights(self):
        u = self.rand()
        rng = f32(len(self.lights))
        index = u32_of(self.mix(f32(0), rng, u))
        return self.lights[min(index, len(self.lights) - 1)]

    def sample_light(self, light):
        u = self.rand()
        v = self.rand()
        return self.a_plus_s_b(self.a_plus_s_b(light["data1"], u, light["data2"]), v, light["data3"])

    def compute_light_pdf(self, intersection):
        shape_intersection = intersection["shape_intersection"]
        intersection_context = intersection["context"]
        light = self.lights[min(shape_intersection["emission_index"], len(self.lights) - 1)]
        light_area = self.length(light["data2"]) * self.length(light["data3"])
        light_area_pdf = f32(1.0) / light_area
        abs_cos_theta = wg_max(f32(0.00001), abs(self.dot(shape_intersection["normal"], v_neg(intersection_context["ray_direction"]))))
        distance = self.length(v_sub(shape_intersection["position"], intersection_context["ray_origin"]))
        distance_squared = distance * distance                                    # pow(distance, 2), section 3
        geometric_term = abs_cos_theta / distance_squared
        light_solid_angle_pdf = light_area_pdf / geometric_term
        number_of_lights = f32(len(self.lights))
        light_selection_pdf = f32(1.0) / number_of_lights
        return light_selection_pdf * light_solid_angle_pdf

    @staticmethod
    def power_heuristic(nf, f_pdf, ng, g_pdf):
        f = f32(nf) * f_pdf
        g = f32(ng) * g_pdf
        return (f * f) / (f * f + g * g)

    def compute_light_radiance(self, intersection, wavelengths):
        shape_intersection = intersection["shape_intersection"]
        context = intersection["context"]
        light = self.sample_lights()
        point_on_light = self.sample_light(light)
        light_dir = self.normalize(v_sub(point_on_light, shape_intersection["position"]))
        shadow_intersection = self.shadow_intersect(dict(origin=shape_intersection["position"], direction=light_dir),
                                                    light["data4"][3], context["index"])
        cos_theta = wg_max(f32(0), self.dot(shape_intersection["normal"], light_dir))
        spec = self.sample_spectrum(light["data4"][0], wavelengths)
        le = v_scale(spec, cos_theta)
        pdf_l = self.compute_light_pdf(shadow_intersection)
        if shadow_intersection["context"]["hit"]:
            pdf_b = cos_theta / PI
            weight_l = self.power_heuristic(1, pdf_l, 1, pdf_b)
            return v_div(v_scale(le, weight_l), pdf_l)
        return (f32(0),) * 4

    # -------------------------------------------------------------- sampling, materials CS:751-774, 814-837
    def cosine_weighted_sample_hemisphere(self, normal):
        u = self.rand()
        v = self.rand()
        r = np.sqrt(u)
        theta = f32(2.0) * PI * v
        sin_t, cos_t = self.m.sincos(theta)
        x = r * cos_t
        y = r * sin_t
        z = np.sqrt(wg_max(f32(0.0), f32(1.0) - u))
        if abs(normal[2]) < f32(0.999):
            up = (f32(0.0), f32(0.0), f32(1.0))
        else:
            up = (f32(1.0), f32(0.0), f32(0.0))
        tangent = self.normalize(self.cross(up, normal))
        bitangent = self.cross(normal, tangent)
        m = self.m
        direction = tuple(m.fma(n, z, m.fma(b, y, t * x)) for t, b, n in zip(tangent, bitangent, normal))
        return direction, z / PI

    def fresnel_s(self, ray_dir, normal, eta1, eta2):
        cosi = wg_min(wg_max(self.dot(ray_dir, normal), f32(-1.0)), f32(1.0))
        eta = eta1 / eta2
        if cosi > 0.0:
            eta = eta2 / eta1
        sint2 = eta * eta * (f32(1.0) - cosi * cosi)
        if sint2 > 1.0:
            return f32(1.0)
        cost = np.sqrt(f32(1.0) - sint2)
        cosi = abs(cosi)
        Rs = ((eta1 * cosi) - (eta2 * cost)) / ((eta1 * cosi) + (eta2 * cost))
        Rp = ((eta2 * cosi) - (eta1 * cost)) / ((eta2 * cosi) + (eta1 * cost))
        return (Rs * Rs + Rp * Rp) / f32(2.0)

    # -------------------------------------------------------------- path_trace CS:119-295
    def path_trace(self, input_ray, wavelengths):
        one4 = (f32(1.0),) * 4
        ray = dict(input_ray)
        depth = 0
        accumulated_radiance = (f32(0.0),) * 4
        beta = one4
        last_bounce_pdf = f32(1.0)
        exclude = MAX_U32_VALUE
        specular_bounce = False
        etaScale = f32(1.0)
        inTransmission = False
        while True:
            self.rays_ext += 1
            intersection = self.intersect(ray, exclude, False)
            shape_intersection = intersection["shape_intersection"]
            intersection_context = intersection["context"]
            if not intersection_context["hit"]:
                self.term["miss"] += 1
                break
            exclude = intersection_context["index"]
            material = shape_intersection["material"]
            if material == LIGHT:
                le = self.sample_spectrum(shape_intersection["emission_index"], wavelengths)
                if depth == 0 or specular_bounce:
                    accumulated_radiance = v_add(accumulated_radiance, v_mul(beta, le))
                else:
                    self.events["mis_light_hit"] += 1
                    pdf_l = self.compute_light_pdf(intersection)
                    weight_b = self.power_heuristic(1, last_bounce_pdf, 1, pdf_l)
                    accumulated_radiance = v_add(accumulated_radiance, v_mul(tuple(weight_b * x for x in le), beta))
                self.term["light"] += 1
                break
            if depth >= self.MAXDEPTH:
                self.term["depth"] += 1
                break
            if inTransmission:
                distance = self.length(v_sub(intersection_context["ray_origin"], shape_intersection["position"]))
                extinction = self.sample_spectrum(self.spectra.shape[0] - 1, wavelengths)
                attenuation = tuple(self.m.exp(-e * distance) for e in extinction)
                self.events["attenuate"] += 1
                beta = v_mul(beta, attenuation)
            if shape_intersection["material"] == DIFFUSE:
                BRDF = v_div(self.sample_spectrum(shape_intersection["reflectance_index"], wavelengths), PI)
                le = self.compute_light_radiance(intersection, wavelengths)
                accumulated_radiance = v_add(accumulated_radiance, v_mul(v_mul(BRDF, le), beta))
                new_direction, last_bounce_pdf = self.cosine_weighted_sample_hemisphere(shape_intersection["normal"])
                cos_theta = abs(self.dot(shape_intersection["normal"], new_direction))
                beta = v_mul(beta, v_div(v_scale(BRDF, cos_theta), last_bounce_pdf))
                ray["origin"] = shape_intersection["position"]
                ray["direction"] = new_direction
                specular_bounce = False
            if shape_intersection["material"] == GLASS:
                eta1 = f32(1.0)
                eta2 = f32(1.5)
                eta = eta1 / eta2
                cos_theta = self.dot(shape_intersection["normal"], ray["direction"])
                reflected = self.fresnel_s(ray["direction"], shape_intersection["normal"], eta1, eta2)
                pr = reflected
                pt = f32(1) - reflected
                u = self.rand()
                current_direction = ray["direction"]
                current_normal = shape_intersection["normal"]
                if cos_theta > 0:
                    eta = f32(1.0) / eta
                    current_normal = v_neg(current_normal)
                if u < pr / (pr + pt):
                    self.events["reflect"] += 1
                    new_direction = self.reflect(current_direction, current_normal)
                    ray["origin"] = shape_intersection["position"]
                    specular_bounce = True
                    exclude = MAX_U32_VALUE
                else:
                    specular_bounce = True
                    exclude = MAX_U32_VALUE
                    self.events["refract"] += 1
                    new_direction = self.refract(current_direction, current_normal, eta)
                    new_direction = self.normalize(new_direction)
                    ray["origin"] = shape_intersection["position"]
                    beta = v_scale(beta, eta * eta)
                    etaScale = etaScale / (eta * eta)
                    inTransmission = not inTransmission
                ray["direction"] = new_direction
            rbeta = v_scale(beta, etaScale)
            max_beta_component = wg_max(rbeta[0], wg_max(rbeta[1], rbeta[2]))
            if depth > 1 and max_beta_component < 1:
                q = wg_max(f32(0), f32(1) - max_beta_component)
                if self.rand() < q:
                    self.term["rr"] += 1
                    break
                beta = v_div(beta, f32(1) - q)
            depth += 1
        self.depth_sum += depth
        return accumulated_radiance

    # -------------------------------------------------------------- colour CS:410-467
    def spectral_to_xyz(self, radiance, wavelengths):
        m = self.m
        integ = f32(106.856895)
        lambda_range = lambda_max - lambda_min
        xyz = []
        for k in range(3):                                                        # row k of transpose(mat3x4(X_BAR, Y_BAR, Z_BAR)) times radiance
            bar = [f32(self.CIE[k][l + 40]) for l in wavelengths]
            xyz.append(m.fma(bar[3], radiance[3], m.fma(bar[2], radiance[2], m.fma(bar[1], radiance[1], bar[0] * radiance[0]))))
        return tuple(x * lambda_range / (integ * f32(4)) for x in xyz)

    def xyz_to_linear_rgb(self, c):
        m = self.m
        r = m.fma(f32(-0.4985314), c[2], m.fma(f32(-1.5371385), c[1], f32(3.2404542) * c[0]))
        g = m.fma(f32(0.0415560), c[2], m.fma(f32(1.8760108), c[1], f32(-0.9692660) * c[0]))
        b = m.fma(f32(1.0572252), c[2], m.fma(f32(-0.2040259), c[1], f32(0.0556434) * c[0]))
        return (r, g, b)

    def tone_map(self, rgb, exposure):
        return tuple(f32(1.0) - self.m.exp(-c * exposure) for c in rgb)

    def gamma_correct(self, rgb):
        r, g, b = rgb
        if r < f32(0.0031308):
            r = r * f32(12.92)
        else:
            r = f32(1.055) * self.m.pow_1_2p4(r) - f32(0.055)
        if g < f32(0.0031308):
            g = g * (f32(12.92) * g)                                              # CS:453 as written: `g *= 12.92 * g`
        else:
            g = f32(1.055) * self.m.pow_1_2p4(g) - f32(0.055)
        if b < f32(0.0031308):
            b = f32(12.92) * b
        else:
            b = f32(1.055) * self.m.pow_1_2p4(b) - f32(0.055)
        return (r, g, b)

    def resolve(self, accumulated_xyz_color, sample: int, exposure=f32(2.2)):
        """CS:109-115 for one pixel: average, XYZ -> RGB, tone map, gamma, textureStore to rgba8unorm."""
        with np.errstate(all="ignore"):
            averaged_xyz_color = v_div(tuple(map(f32, accumulated_xyz_color)), f32(sample))
            rgb_color = self.xyz_to_linear_rgb(averaged_xyz_color)
            rgb_color_ldr = self.gamma_correct(self.tone_map(rgb_color, f32(exposure)))
            return tuple(unorm8(c) for c in rgb_color_ldr) + (255,)

    # -------------------------------------------------------------- camera CS:469-500
    def camera_basis(self):
        cam = self.camera
        w = self.normalize(v_sub(cam["eye"], cam["lookat"]))
        u = self.normalize(self.cross(cam["up"], w))
        v = self.cross(w, u)
        return u, v, w

    def camera_film_parameters(self, screen_pos, screen_size):
        G = f32(self.GRID_SIZE)
        s = (f32(screen_pos[0]) + (f32(self.sample % self.GRID_SIZE) + self.rand()) / G) / f32(screen_size[0])
        t = (f32(screen_size[1]) - f32(screen_pos[1]) + (f32(self.sample % self.GRID_SIZE) + self.rand()) / G) / f32(screen_size[1])
        return s, t

    def camera_ray(self, screen_pos, screen_size):
        cam = self.camera
        aspect_ratio = cam["viewport_width"] / cam["viewport_height"]
        viewport_height = f32(2.0) * f32(_libm.tan(float(cam["focal_length"] / f32(2.0))))   # tan: double libm, rounded once (section 3)
        viewport_width = aspect_ratio * viewport_height
        u, v, w = self.camera_basis()
        horizontal = v_scale(u, viewport_width)
        vertical = v_scale(v, viewport_height)
        lower_left_corner = v_sub(v_sub(v_sub(cam["eye"], v_div(horizontal, f32(2.0))), v_div(vertical, f32(2.0))), w)
        s, t = self.camera_film_parameters(screen_pos, screen_size)
        film = self.a_plus_s_b(self.a_plus_s_b(lower_left_corner, s, horizontal), t, vertical)
        return dict(origin=cam["eye"], direction=self.normalize(v_sub(film, cam["eye"])))

    # -------------------------------------------------------------- main CS:77-117
    def reset_statistics(self):
        self.n_rand = self.rays_ext = self.rays_shadow = self.prim_tests = self.depth_sum = self.paths = 0
        self.term = dict(miss=0, light=0, depth=0, rr=0)
        self.events = dict(reflect=0, refract=0, attenuate=0, mis_light_hit=0)

    def main(self, gid_x: int, gid_y: int, sample: int, t_min=f32(0.001)):
        """One invocation; returns the XYZ this sample adds to alt_color_buffer[pixel] (None outside the screen)."""
        self.sample, self.t_min = sample, f32(t_min)
        cam = self.camera
        screen_size = (u32_of(cam["viewport_width"]), u32_of(cam["viewport_height"]))
        screen_pos = (gid_x, gid_y)
        if gid_x >= screen_size[0] or gid_y >= screen_size[1]:
            return None
        self.paths += 1
        self.seed = [screen_pos[1], (screen_pos[0] * 100) & M32, sample & M32, self.tea(screen_pos[0], (screen_pos[1] * 100) & M32)]
        with np.errstate(all="ignore"):
            ray = self.camera_ray(screen_pos, screen_size)
            wavelengths = self.sample_wavelengths()
            radiance = self.path_trace(ray, wavelengths)
            return self.spectral_to_xyz(radiance, wavelengths)


def unorm8(c) -> int:
    """rgba8unorm store (section 3): NaN -> 0, clamp to [0, 1], round half up."""
    if c != c:
        return 0
    c = wg_min(wg_max(f32(c), f32(0.0)), f32(1.0))
    return int(np.floor(c * f32(255.0) + f32(0.5)))


def wg_max(a, b):
    """max() as the C library and the GPU evaluate it: a NaN operand loses."""
    if a != a:
        return b
    if b != b:
        return a
    return b if b > a else a


def wg_min(a, b):
    if a != a:
        return b
    if b != b:
        return a
    return b if b < a else a

// sceneLoader.js -- scene JSON -> the byte buffers the compute shader binds (Node.js host side).
//
// The reference's src/sceneLoader.js is a 5-line stub; its packing code lives inline in
// src/main.js:114-393.  This module is that contract as one function, producing the buffers the
// N-API addon (crt_addon.c) hands to libcrt by binding number.  It is the JavaScript twin of
// computeraytracer_b200/scene_loader.py (the tested packer: node is absent from the build image, so this
// file ships as source only -- see INTEGRATION.md).
//
//   binding 3  planar patches   64-byte stride   main.js:139-209
//   binding 4  camera           16 x f32         main.js:313-332
//   binding 5  CIE X|Y|Z        3 x 471 f32      main.js:380-393
//   binding 6  spectra          n x 301 f32      main.js:334-378
//   binding 7  lights           80-byte stride   main.js:255-296
//   binding 8  primitives       80-byte stride   main.js:114-137, 211-246
'use strict';

const LAMBDA_MIN = 400;
const LAMBDA_MAX = 700;
const N_LAMBDA = LAMBDA_MAX - LAMBDA_MIN + 1;
const TYPE_INDEX = { diffuse: 0, light: 1, glass: 2 };
const CATEGORY = { patch: 0, sphere: 1, triangle: 2 };   // 2 = libcrt's triangle extension
const PRIM_STRIDE = 80;
const PATCH_STRIDE = 64;

// piecewise-linear lookup, clamped at both ends (main.js:340-356), evaluated in double
function sampleSpectrum(spectrum, lambda) {
  const wl = spectrum.wavelength;
  const val = spectrum.value;
  let index = wl.findIndex((e) => e >= lambda);
  let i0 = Math.max(index - 1, 0);
  let i1 = Math.min(index, wl.length - 1);
  if (index < 0) { i0 = i1 = wl.length - 1; }
  if (wl[i0] === wl[i1]) return val[i0];
  const t = (lambda - wl[i0]) / (wl[i1] - wl[i0]);
  return val[i0] * (1 - t) + val[i1] * t;
}

function writeVec3(f32, byteOffset, v) {
  const o = byteOffset >> 2;
  f32[o] = v[0]; f32[o + 1] = v[1]; f32[o + 2] = v[2];
}

// One 80-byte Primitive record (ComputeShader.wgsl:41-47)
function writePrimitive(f32, u32, slot, category, d1, d2, d3, emission, reflectance, type, index) {
  const base = slot * PRIM_STRIDE;
  u32[base >> 2] = category;
  writeVec3(f32, base + 16, d1);
  writeVec3(f32, base + 32, d2);
  writeVec3(f32, base + 48, d3);
  const q = (base + 64) >> 2;
  u32[q] = emission; u32[q + 1] = reflectance; u32[q + 2] = type; u32[q + 3] = index;
}

/**
 * @param scene      object with the schema of src/scenes/cornell.json
 * @param cie        object with CIE_X, CIE_Y, CIE_Z (src/scenes/CIE.json)
 * @param triangles  optional Float32Array of n*9 floats (v0, v1, v2 per triangle): category-2 extension
 * @returns {width, height, bindings: {3..8: ArrayBuffer}, spectrumIndex}
 */
function loadScene(scene, cie, triangles = null, triangleReflectance = 'white') {
  const keyIndex = {};
  Object.keys(scene.spectra).forEach((name, i) => { keyIndex[name] = i; });   // main.js:161-164
  const patches = scene.objects.patches || [];
  const spheres = scene.objects.spheres || [];
  const nTri = triangles ? triangles.length / 9 : 0;
  const nPrims = patches.length + spheres.length + nTri;

  const primBuf = new ArrayBuffer(nPrims * PRIM_STRIDE);
  const pf = new Float32Array(primBuf);
  const pu = new Uint32Array(primBuf);
  const patchBuf = new ArrayBuffer(patches.length * PATCH_STRIDE);
  const qf = new Float32Array(patchBuf);
  const qu = new Uint32Array(patchBuf);
  const lightSlots = [];

  patches.forEach((p, i) => {
    const e = keyIndex[p.emission], r = keyIndex[p.reflectance], t = TYPE_INDEX[p.type];
    const b = i * PATCH_STRIDE;                                     // main.js:172-209
    writeVec3(qf, b, p.origin); writeVec3(qf, b + 16, p.edge1); writeVec3(qf, b + 32, p.edge2);
    qu[(b + 44) >> 2] = e; qu[(b + 48) >> 2] = r; qu[(b + 52) >> 2] = t; qu[(b + 56) >> 2] = i;
    writePrimitive(pf, pu, i, CATEGORY.patch, p.origin, p.edge1, p.edge2, e, r, t, i);   // main.js:224-233
    if (p.type === 'light') lightSlots.push(i);
  });
  spheres.forEach((s, k) => {
    const i = patches.length + k;                                   // main.js:234-244
    const rr = [s.radius, s.radius, s.radius];
    writePrimitive(pf, pu, i, CATEGORY.sphere, s.center, rr, [0, 0, 0], keyIndex[s.emission], keyIndex[s.reflectance],
      TYPE_INDEX[s.type], i);
    if (s.type === 'light') lightSlots.push(i);
  });
  for (let k = 0; k < nTri; k++) {
    const i = patches.length + spheres.length + k;
    const v = triangles.subarray(9 * k, 9 * k + 9);
    const f = Math.fround;
    const e1 = [f(v[3] - v[0]), f(v[4] - v[1]), f(v[5] - v[2])];
    const e2 = [f(v[6] - v[0]), f(v[7] - v[1]), f(v[8] - v[2])];
    writePrimitive(pf, pu, i, CATEGORY.triangle, [v[0], v[1], v[2]], e1, e2, keyIndex.dark || 0,
      keyIndex[triangleReflectance], TYPE_INDEX.diffuse, i);
  }

  // lights = the primitives whose type is "light", same 80-byte records (main.js:255-296)
  const lightBuf = new ArrayBuffer(lightSlots.length * PRIM_STRIDE);
  const lb = new Uint8Array(lightBuf);
  const pb = new Uint8Array(primBuf);
  lightSlots.forEach((slot, k) => { lb.set(pb.subarray(slot * PRIM_STRIDE, (slot + 1) * PRIM_STRIDE), k * PRIM_STRIDE); });

  const cam = scene.camera;                                         // main.js:313-324
  const camera = new Float32Array([...cam.eye, 0, ...cam.lookat, 0, ...cam.up, cam.width, cam.height, cam.focalLength, 0, 0]);

  const names = Object.keys(scene.spectra);                         // main.js:358-378
  const spectra = new Float32Array(names.length * N_LAMBDA);
  names.forEach((name, s) => {
    for (let i = 0; i < N_LAMBDA; i++) spectra[s * N_LAMBDA + i] = sampleSpectrum(scene.spectra[name], LAMBDA_MIN + i);
  });
  const cieTable = new Float32Array([...cie.CIE_X, ...cie.CIE_Y, ...cie.CIE_Z]);   // main.js:380-393

  return {
    width: cam.width,
    height: cam.height,
    spectrumIndex: keyIndex,
    bindings: { 3: patchBuf, 4: camera.buffer, 5: cieTable.buffer, 6: spectra.buffer, 7: lightBuf, 8: primBuf },
  };
}

module.exports = { loadScene, sampleSpectrum, LAMBDA_MIN, LAMBDA_MAX, N_LAMBDA, TYPE_INDEX, CATEGORY };

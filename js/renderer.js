// renderer.js -- Main()/frame() of the reference's src/main.js against libcrt instead of WebGPU.
//
// SOURCE ONLY here (no node in the build image); the same sequence is tested through
// computeraytracer_b200/renderer.py.  Usage:
//
//   const { Renderer } = require('./renderer');
//   const r = new Renderer(0);                       // main.js:8-9
//   r.setScene(require('./scene.json'), require('./CIE.json'));   // main.js:114-393
//   r.frame(64);                                     // 64 x frame(), main.js:584-621
//   const rgba = r.framebuffer();                    // Uint8Array W*H*4, top-left origin (binding 0)
'use strict';

const addon = require('./build/Release/crt_addon.node');
const { loadScene } = require('./sceneLoader');

class Renderer {
  constructor(device = 0) {
    this.h = addon.create(device);   // throws when there is no B200: libcrt has no CPU fallback
    this.width = 0;
    this.height = 0;
  }

  setScene(scene, cie, triangles = null) {
    const packed = loadScene(scene, cie, triangles);
    for (const [binding, buf] of Object.entries(packed.bindings)) addon.setBinding(this.h, Number(binding), buf);
    this.width = packed.width;
    this.height = packed.height;
    return packed;
  }

  setParams(params) { addon.setParams(this.h, params); }      // {maxDepth, gridSize, tMin, exposure}
  reset() { addon.reset(this.h); }
  frame(n = 1) { addon.dispatchFrames(this.h, n); }            // sample++ ; one sample per pixel -- n times
  get sample() { return new Uint32Array(addon.readBinding(this.h, 2))[0]; }
  accumulator() { return new Float32Array(addon.readBinding(this.h, 1)); }   // XYZ sums, 16-byte stride
  framebuffer() { return new Uint8Array(addon.readBinding(this.h, 0)); }     // RGBA8
  counters() { return addon.counters(this.h); }
  close() { addon.destroy(this.h); }
}

module.exports = { Renderer };

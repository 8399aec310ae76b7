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
  // device: one GPU index, or an array of indices = one handle over several GPUs of this process (crt_create_multi;
  // shardMode 1 = tiles, 2 = split-spp): frames are sharded, readbacks return the combined image of the first device
  constructor(device = 0, shardMode = 2) {
    this.h = Array.isArray(device) ? addon.createMulti(device, shardMode) : addon.create(device);   // throws when there is no B200: libcrt has no CPU fallback
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
  // one process per GPU instead: rank 0 calls Renderer.commUniqueId(), sends the 128 bytes to the other ranks (any channel),
  // every rank calls commInit; combine(dst) is then the collective that leaves the combined image on rank dst
  static commUniqueId() { return addon.commUniqueId(); }
  commInit(id, rank, world, shardMode = 2) { addon.commInit(this.h, id, rank, world, shardMode); }
  combine(dst = 0) { addon.combine(this.h, dst); }
  close() { addon.destroy(this.h); }                           // frees the device memory now
}

module.exports = { Renderer };

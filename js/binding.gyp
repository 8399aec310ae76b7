{
  "targets": [{
    "target_name": "crt_addon",
    "sources": ["crt_addon.c"],
    "include_dirs": ["../include"],
    "libraries": ["-L<(module_root_dir)/../computeraytracer_b200", "-lcrt",
                  "-Wl,-rpath,<(module_root_dir)/../computeraytracer_b200"],
    "defines": ["NAPI_VERSION=8"]
  }]
}

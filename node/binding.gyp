{
  "targets": [{
    "target_name": "brotligpu",
    "sources": ["addon.cc"],
    "include_dirs": ["../include"],
    "libraries": ["-L<(module_root_dir)/../brotli_lib_b200", "-lbrotligpu", "-Wl,-rpath,<(module_root_dir)/../brotli_lib_b200"],
    "cflags_cc": ["-std=c++17"]
  }]
}

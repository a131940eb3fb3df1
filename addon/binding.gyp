{
  "targets": [{
    "target_name": "beholder_napi",
    "sources": ["beholder_napi.c"],
    "include_dirs": ["../include"],
    "libraries": ["-L<(module_root_dir)/../beholder_b200", "-lbeholder_gpu", "-Wl,-rpath,<(module_root_dir)/../beholder_b200"]
  }]
}

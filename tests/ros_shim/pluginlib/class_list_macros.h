// TEST INFRASTRUCTURE ONLY. Stand-in for <pluginlib/class_list_macros.h>: registration is a no-op here.
#pragma once
#define PLUGINLIB_EXPORT_CLASS(class_type, base_class_type)

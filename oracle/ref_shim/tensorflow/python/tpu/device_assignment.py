class DeviceAssignment:
    pass

ssh test@127.0.0.1 "kill -9 2160"
ssh test@127.0.0.1 "kill -9 2161"
ssh test@127.0.0.1 "kill -9 2162"
kill -9 2109
